#!/usr/bin/env python
"""Extract the reference's golden HDF5 outputs into small fixtures (tests/golden/*.npz).

Run in the build container, where /root/reference exists:

    python tests/golden/make_golden.py

The GPU box has no /root/reference, so tests read only the committed .npz files.  The source
files are DLBC's own regression data, tests/runnable/tests/**/reference-data/*.h5 (dataset
/OutArray), read with dlbc_b200/h5lite.py.  Values are stored bit-exactly (see codec.py).
"""
import glob
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import codec  # noqa: E402
from dlbc_b200.h5lite import read_dataset  # noqa: E402

T = "/root/reference/tests/runnable/tests/"


def g(pattern):
    files = glob.glob(T + pattern)
    assert len(files) == 1, (pattern, files)
    return read_dataset(files[0])


def main():
    # --- lb/merging-droplets-3d (BASELINE config 1): densities + colour at t = 0, 10, 1000
    out = {}
    for t in (0, 10, 1000):
        for name in ("density-red", "density-blue", "colour-red-blue"):
            a = g(f"lb/merging-droplets-3d/reference-data/{name}-*t{t:08d}.h5")
            out[f"{name}-t{t}"] = a
        # colour is exactly red - blue in the reference's files: keep only the densities
        assert np.array_equal(out[f"colour-red-blue-t{t}"], out[f"density-red-t{t}"] - out[f"density-blue-t{t}"])
        del out[f"colour-red-blue-t{t}"]
    codec.save(os.path.join(HERE, "merging-droplets-3d.npz"), **out)

    # --- lb/bdist2-sc-3d: full populations.  The system is lamellae along z, uniform in x and y;
    # check that on the golden data and keep one z-column [0, 0, :, :].
    out = {}
    for t in (0, 10, 100):
        for c in ("red", "blue"):
            p = g(f"lb/bdist2-sc-3d/reference-data/population-{c}-*t{t:08d}.h5")
            assert np.array_equal(p, np.broadcast_to(p[0:1, 0:1], p.shape)), "not uniform in x,y"
            out[f"population-{c}-t{t}"] = p[0, 0]
            d = g(f"lb/bdist2-sc-3d/reference-data/density-{c}-*t{t:08d}.h5")
            assert np.array_equal(d, np.broadcast_to(d[0:1, 0:1], d.shape))
            out[f"density-{c}-t{t}"] = d[0, 0]
        col = g(f"lb/bdist2-sc-3d/reference-data/colour-red-blue-*t{t:08d}.h5")
        out[f"colour-red-blue-t{t}"] = col[0, 0]
    codec.save(os.path.join(HERE, "bdist2-sc-3d.npz"), **out)

    # --- lb/bdist2-3d: 4^3, no Shan-Chen
    out = {}
    for t in (0, 10):
        for c in ("red", "blue"):
            out[f"population-{c}-t{t}"] = g(f"lb/bdist2-3d/reference-data/population-{c}-*t{t:08d}.h5")
            out[f"density-{c}-t{t}"] = g(f"lb/bdist2-3d/reference-data/density-{c}-*t{t:08d}.h5")
    codec.save(os.path.join(HERE, "bdist2-3d.npz"), **out)

    # --- technical/checkpoint-3d: its parameter file exists only as the `input` attribute of the output files
    # (the test's input file IS reference-data/colour-red-blue-checkpoint-3d-t00000000.h5): extracted to in/
    from dlbc_b200 import h5lite as _h5
    _lines = _h5.read_attributes(os.path.join(T, "technical/checkpoint-3d/reference-data/colour-red-blue-checkpoint-3d-t00000000.h5"))["input"]
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "in", "checkpoint-3d.in"), "w") as _f:
        _f.write("\n".join(_lines) + "\n")
    # --- technical/checkpoint-3d: restart KAT (checkpoint populations + mask -> 5 steps)
    out = {}
    R = "technical/checkpoint-3d/reference-data/"
    for c, name in enumerate(("red", "blue")):
        out[f"cp-fluids{c}-t0"] = g(R + f"cp/cp-fluids{c}-checkpoint-3d-t00000000.h5")
        out[f"cp-fluids{c}-t5"] = g(R + f"cp/cp-fluids{c}-checkpoint-3d-t00000005.h5")
        for t in (0, 5):
            out[f"population-{name}-t{t}"] = g(R + f"population-{name}-checkpoint-3d-t{t:08d}.h5")
            out[f"density-{name}-t{t}"] = g(R + f"density-{name}-checkpoint-3d-t{t:08d}.h5")
    for t in (0, 5):
        out[f"cp-mask-t{t}"] = g(R + f"cp/cp-mask-checkpoint-3d-t{t:08d}.h5")
        out[f"colour-red-blue-t{t}"] = g(R + f"colour-red-blue-checkpoint-3d-t{t:08d}.h5")
    codec.save(os.path.join(HERE, "checkpoint-3d.npz"), **out)

    # --- lb/laplace-droplet-3d: SphereFrac init; t = 0, 10, 100 densities
    out = {}
    for t in (0, 10, 100):
        for c in ("red", "blue"):
            out[f"density-{c}-t{t}"] = g(f"lb/laplace-droplet-3d/reference-data/density-{c}-*t{t:08d}.h5")
    codec.save(os.path.join(HERE, "laplace-droplet-3d.npz"), **out)

    # --- lb/poiseuille-2d (D2Q9): bounce-back, SecondOrder, globalAcc, velocity/force dumps
    out = {}
    for t in (0, 1000, 100000):
        out[f"mask-t{t}"] = g(f"lb/poiseuille-2d/reference-data/mask-*t{t:08d}.h5")
        for c in ("red", "blue"):
            for f in ("density", "velocity", "force"):
                out[f"{f}-{c}-t{t}"] = g(f"lb/poiseuille-2d/reference-data/{f}-{c}-*t{t:08d}.h5")
    codec.save(os.path.join(HERE, "poiseuille-2d.npz"), **out)
    for ext in (".in", ".json"):
        shutil.copy(T + "lb/poiseuille-2d/poiseuille-2d" + ext, os.path.join(HERE, "in", "poiseuille-2d" + ext))

    # --- lb/init-2d: the random initialisers (np = 4, nc = [2, 2]) pin Mt19937 + uniform
    out = {}
    for name in ("eqdistperturb", "eqdistperturbfrac", "eqdistrandom", "constrandom"):
        for c in ("red", "blue"):
            out[f"{name}-{c}"] = g(f"lb/init-2d/reference-data/population-{c}-fluid-init-{name}-2d-t00000000.h5")
    codec.save(os.path.join(HERE, "init-2d-random.npz"), **out)

    # --- the 2-D (D2Q9) spec cases: what their fast runs compare (t = 0 and the fast run's last dump), whole arrays;
    # the parameter files and JSON specs are copied to in/ byte for byte
    import shutil
    # (+ the last dump of the FULL runs of the two Shan-Chen cases: t = 100 and t = 1000)
    for case, times, data in (("bdist2-2d", (0, 10), ("density-red", "density-blue", "population-red", "population-blue")),
                              ("bdist2-sc-2d", (0, 10, 100), ("colour-red-blue", "density-red", "density-blue", "population-red",
                                                              "population-blue")),
                              ("merging-droplets-2d", (0, 100, 1000), ("colour-red-blue", "density-red", "density-blue",
                                                                       "population-red", "population-blue")),
                              ("laplace-droplet-2d", (0, 100), ("colour-red-blue", "density-red", "density-blue"))):
        out = {}
        for t in times:
            for d in data:
                out[f"{d}-t{t}"] = g(f"lb/{case}/reference-data/{d}-{case}*-t{t:08d}.h5")  # (laplace: ...-gcc3.6-...)
        codec.save(os.path.join(HERE, case + ".npz"), **out)
        for ext in (".in", ".json"):
            shutil.copy(T + f"lb/{case}/{case}{ext}", os.path.join(HERE, "in", case + ext))

    # --- technical/checkpoint-2d (d2q9): input file an .h5 (extracted to in/checkpoint-2d.in), restart from cp files, t = 0, 10
    C2 = "technical/checkpoint-2d/"
    tid = "20150321T122217"
    lines2 = _h5.read_attributes(T + C2 + f"reference-data/colour-red-blue-checkpoint-2d-{tid}-t00000000.h5")["input"]
    with open(os.path.join(HERE, "in", "checkpoint-2d.in"), "w") as _f:
        _f.write("\n".join(lines2) + "\n")
    shutil.copy(T + C2 + "checkpoint-2d.json", os.path.join(HERE, "in", "checkpoint-2d.json"))
    out = {}
    for name in ("cp-fluids0", "cp-fluids1", "cp-mask"):
        out[f"{name}-t0"] = g(C2 + f"reference-data/cp/{name}-checkpoint-2d-{tid}-t00000000.h5")
    for t in (0, 10):
        for d in ("colour-red-blue", "density-red", "density-blue", "population-red", "population-blue"):
            out[f"{d}-t{t}"] = g(C2 + f"reference-data/{d}-checkpoint-2d-{tid}-t{t:08d}.h5")
    codec.save(os.path.join(HERE, "checkpoint-2d.npz"), **out)

    # --- lb/init-2d: 32 parameter files (np = 4, nc = [2, 2]), one per fluid / mask initialiser, each with its t = 0
    # dumps.  Kept as SHA-256 of the arrays (NaN canonicalised: FluidInit.None leaves double.init in the field) plus the
    # three arrays the two File cases read back in; the parameter files are copied to in/init-2d/ byte for byte.
    import hashlib
    import json
    import re
    os.makedirs(os.path.join(HERE, "in", "init-2d"), exist_ok=True)
    hashes = {}
    for inp in sorted(glob.glob(T + "lb/init-2d/*.in")):
        case = os.path.basename(inp)[:-3]
        shutil.copy(inp, os.path.join(HERE, "in", "init-2d", case + ".in"))
        hashes[case] = {}
        for f in sorted(glob.glob(T + f"lb/init-2d/reference-data/*-{case}-t00000000.h5")):
            stem = re.sub(r"-\d{8}T\d{6}", "", os.path.basename(f)[:-3])
            hashes[case][stem] = codec.digest(read_dataset(f))
    with open(os.path.join(HERE, "init-2d.json"), "w") as fh:
        json.dump(hashes, fh, indent=1, sort_keys=True)
    R = "lb/init-2d/reference-data/"
    codec.save(os.path.join(HERE, "init-2d-files.npz"),
               **{"population-red-fluid-init-file-2d-t00000000": g(R + "population-red-fluid-init-file-2d-t00000000.h5"),
                  "population-blue-fluid-init-file-2d-t00000000": g(R + "population-blue-fluid-init-file-2d-t00000000.h5"),
                  "mask-mask-init-file-2d-t00000000": g(R + "mask-mask-init-file-2d-t00000000.h5")})

    for f in sorted(glob.glob(os.path.join(HERE, "*.npz"))):
        print(f"{os.path.getsize(f):9d}  {os.path.basename(f)}")


if __name__ == "__main__":
    main()
