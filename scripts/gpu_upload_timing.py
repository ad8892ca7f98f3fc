import sys, time, numpy as np
sys.path.insert(0, '.')
import bench
from dlbc_b200 import cases, init as I, lib as dlib
from dlbc_b200.lattice import Lattice
G = Lattice((512,512,512), 2, **cases.SC2)
lib = dlib.load()
views=[]
for k in range(2):
    v,p = bench.pinned_array(lib, (4,512,512,19)); v[...] = 0.37/19; views.append(v)
for rep in range(3):
    G.sync(); t0=time.perf_counter(); n=bench.fill_lattice(G, (512,512,512), 2, views); t1=time.perf_counter(); G.sync(); t2=time.perf_counter()
    print("fill returned %.3f s, synced %.3f s, %.1f GB/s" % (t1-t0, t2-t0, n/(t2-t0)/1e9))
