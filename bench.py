#!/usr/bin/env python
"""bench.py -- MLUPS of the DLBC timestep (D3Q19, two-component Shan-Chen) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout (rank 0).  A "step" is one lattice-Boltzmann timestep (the statements of
runTimeloop, reference src/main.d:179-195) over the whole lattice.

* workload (N = 1): BASELINE.json configs[2], D3Q19 two-component Shan-Chen spinodal decomposition
  at 512^3 (the configuration the metric is quoted on); N > 1: weak scaling, 512^3 per GPU as
  z-slabs of a 512 x 512 x (512 N) lattice, ghost planes exchanged with NCCL send/recv.
* value: MLUPS = 1e-6 * nx*ny*nz * K / t (reference definition, src/dlbc/timers.d:222-230), state
  resident in HBM, t from CUDA events on the compute stream, max over ranks.
* e2e: the same metric for the whole job a host program runs through the C ABI with HOST buffers:
  upload of the initial populations from pinned host memory + K steps + download of the density
  fields (what dumpLBData writes), all inside the timed region.
* roofline: algorithmic bytes per site update (2*C*19*8 + 1 = 609 B at C = 2) over the dominant
  kernel's CUDA-event duration, against MEASURED_PEAKS.json's copy bandwidth.
* cpu_baseline / --impl reference: the CPU restatement of the reference timestep (oracle/, "port":
  the D reference cannot be built in this image) on the box's host cores, bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "MLUPS (D3Q19 2-comp Shan-Chen)"
B_ALG = {1: 2 * 1 * 19 * 8 + 1, 2: 2 * 2 * 19 * 8 + 1}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return None
        time.sleep(0.05)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            return None
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "samples": len(sm), "reasons": sorted(reasons)}


# ---- CPU arm: the oracle port on the host cores ---------------------------------------------------
def cpu_port_mlups(sample_n, steps, warmup, ncomp=2):
    """The reference's CPU algorithm (oracle/ restatement) on a sample_n^3 sample of the workload."""
    from dlbc_b200 import cases
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    params, pops, _ = cases.spinodal(sample_n)
    p = dict(params)
    gn, nc = p.pop("gn"), p.pop("components")
    L = O.OracleLattice(gn, nc, **p, nthreads=cores)
    for c, a in enumerate(pops):
        L.set_populations(c, a)
    L.exchange_halo()
    L.step(max(1, warmup))
    t0 = time.perf_counter()
    L.step(steps)
    dt = time.perf_counter() - t0
    return sample_n ** 3 * steps / dt * 1e-6, dt, cores


def workload_config(args, world, gn, gl, ncomp, porous):
    """The `config` object both arms print (the reference arm times a bounded SAMPLE of this workload)."""
    sites_local = gl[0] * gl[1] * gl[2]
    return {"workload": ("D3Q19 two-fluid Shan-Chen flow through a random porous bounce-back geometry, "
                         "30 % solid (BASELINE configs[3])" if porous else
                         "D3Q19 two-component Shan-Chen spinodal decomposition (BASELINE configs[2])"
                         if ncomp == 2 else "D3Q19 single-fluid BGK periodic (BASELINE configs[1])"),
            "lattice": list(gn), "components": ncomp, "decomposition": f"{world} z-slab(s) of {list(gl)}",
            "arithmetic": args.arith, "kernel": args.kernel,
            "l2": "inputs exceed L2 (populations are %.1f GB per GPU)" % (2 * ncomp * 19 * 8 * sites_local / 1e9)}


def bench_lattice(args, world):
    n = args.size
    strong = args.scaling == "strong"
    gn = (n, n, n if strong else n * world)   # weak: `size`^3 per GPU stacked in z; strong: `size`^3 in total
    if args.shape:
        sx, sy, sz = (int(v) for v in args.shape.split(","))
        gn = (sx, sy, sz if strong else sz * world)
    return gn, (gn[0], gn[1], gn[2] // world)


def cpu_config1_mlups(steps):
    """BASELINE configs[0] as the reference names it: merging-droplets-3d (64^3) on 4 ranks -> 4 oracle threads."""
    from dlbc_b200 import cases
    from oracle import oracle as O
    params, pops, _ = cases.merging_droplets(64)
    p = dict(params)
    gn, nc = p.pop("gn"), p.pop("components")
    L = O.OracleLattice(gn, nc, **p, nthreads=4)
    for c, a in enumerate(pops):
        L.set_populations(c, a)
    L.exchange_halo()
    L.step(3)
    t0 = time.perf_counter()
    L.step(steps)
    dt = time.perf_counter() - t0
    return 64 ** 3 * steps / dt * 1e-6, dt


def run_reference(args, rank):
    """The reference's CPU algorithm (oracle port: the D reference cannot be built here) on all host cores, same
    metric / config / steps / warm-up as our arm; each step runs on a bounded sample of the workload."""
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = args.cpu_sample
    steps, warmup = max(args.steps, 1), max(args.warmup, 0)
    # size the sample so that warm-up + timed steps stay within ~2 minutes: calibrate on 2 steps
    mlups, dt, cores = cpu_port_mlups(n, 2, 1)
    per_step = dt / 2
    budget = 120.0
    if (steps + warmup) * per_step > budget:
        n = max(32, int(n * (budget / ((steps + warmup) * per_step)) ** (1 / 3)) // 16 * 16)
    mlups, dt, cores = cpu_port_mlups(n, steps, warmup)
    gn, gl = bench_lattice(args, world)
    line = {
        "impl": "reference", "metric": METRIC, "value": mlups, "unit": "MLUPS", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, world, gn, gl, args.components, args.workload == "porous"),
        "cpu_baseline": {"value": mlups, "unit": "MLUPS", "cores": cores, "kind": "port",
                         "sample": f"{n}^3 sample of the workload, {steps} timesteps after {warmup} warm-up steps, {dt:.1f} s; "
                                   "CPU restatement of the DLBC timestep (oracle/, OpenMP x-slabs standing in for MPI ranks): "
                                   "the D reference needs dmd/ldc/gdc + MPICH + parallel HDF5, none of which exist in this image"},
        "e2e": {"value": mlups, "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- GPU arm -----------------------------------------------------------------------------------------
def fill_lattice(G, gn_local, ncomp, pinned_views, timed=False):
    """Upload a perturbed-equilibrium state layer block by layer block through the C ABI."""
    nx = gn_local[0]
    xc = pinned_views[0].shape[0]
    nbytes = 0
    for c in range(ncomp):
        for i, x0 in enumerate(range(0, nx, xc)):
            cnt = min(xc, nx - x0)
            blk = pinned_views[(c + i) % len(pinned_views)]
            G.set_populations_xrange(c, blk[:cnt], x0)
            nbytes += blk[:cnt].nbytes
    return nbytes


def pinned_array(lib, shape):
    import ctypes as C
    n = int(np.prod(shape)) * 8
    ptr = C.c_void_p()
    rc = lib.dlbc_cuda_host_alloc(C.byref(ptr), n)
    if rc:
        raise RuntimeError("dlbc_cuda_host_alloc failed")
    buf = (C.c_double * (n // 8)).from_address(ptr.value)
    return np.frombuffer(buf, dtype=np.float64).reshape(shape), ptr


def _bitwise_equal(a, b):
    return a.shape == b.shape and bool((a.view(np.uint64) == b.view(np.uint64)).all())


def parity_check(args, device):
    """Driver-visible parity evidence for the kernel the bench times: on a 512 x 512 x 64 lattice (the band /
    segment / lag geometry of the single-launch kernels depends on nx, ny only, so this IS the 512^3 regime:
    16 bands of 32 rows, 8 (TMA) or 4 (wave) x-segments per row, lag 4, every SM busy) step the SAME uploaded
    state a few times with the benchmarked kernel, the register-path wave kernel and the two-kernel schedule
    and compare the downloaded populations bit for bit (same site functions => identical bits)."""
    from dlbc_b200 import cases, init as I
    from dlbc_b200.lattice import Lattice
    gn = (args.size, args.size, 64) if not args.shape else tuple(int(v) for v in args.shape.split(","))[:2] + (64,)
    if args.components != 2:
        return None
    steps, xc = 4, 32
    layers = [cases.synthetic_layers(gn, xc, 0.37, 0.02, I.BDIST2, 1000 + k) for k in range(3)]
    kernels = [args.kernel, "wave", "two_pass"] if args.kernel != "two_pass" else ["two_pass", "wave"]
    ctxs, names = [], []
    try:
        for k in kernels:
            G = Lattice(gn, 2, **cases.SC2, arithmetic=args.arith, kernel=k, device=device)
            for c in range(2):
                for i, x0 in enumerate(range(0, gn[0], xc)):
                    G.set_populations_xrange(c, layers[(c + i) % 3][: min(xc, gn[0] - x0)], x0)
            names.append(G.kernel_variant)
            G.step(steps)
            ctxs.append(G)
        same, checked = True, 0
        for c in range(2):
            for x0 in range(0, gn[0], 64):
                cnt = min(64, gn[0] - x0)
                ref = ctxs[-1].get_populations_xrange(c, x0, cnt)
                for G in ctxs[:-1]:
                    same = same and _bitwise_equal(G.get_populations_xrange(c, x0, cnt), ref)
                checked += ref.size
        finite = bool(np.isfinite(ref).all())
    finally:
        for G in ctxs:
            G.close()
    return {"bitexact": bool(same and finite), "kernels": names, "lattice": list(gn), "steps": steps,
            "arithmetic": args.arith, "values_compared": int(checked)}


def decomposition_check(args, dist, rank, world, local_rank, dlib):
    """Driver-visible multi-GPU correctness (the reference runs every test on several process grids against one
    golden, tests/runnable/dlbct/run.py:28-30): the same case on the N-rank NCCL communicator and on ONE rank
    must give bit-identical populations (EXACT arithmetic; the per-site arithmetic does not depend on the
    split).  Two cases: thin slabs (AUTO -> two-kernel schedule) and slabs of 160 planes (AUTO -> the
    single-launch kernel, forced into several bands)."""
    import torch
    from dlbc_b200 import slabs
    from dlbc_b200.lattice import Lattice
    from dlbc_b200 import cases, init as I
    out = {}
    for name, planes, opts in (("thin_slabs", 32, {}), ("thick_slabs", 160, {"wave_tyb": 12})):
        gn = (64, 48, planes * world)
        rngs = [np.random.default_rng(77 + c) for c in range(2)]
        pops = [(r0 + 0.05 * g.random(gn))[..., None] * I.eq_dist_unity(cases.d3q19, I.BDIST2)
                for r0, g in zip((0.6, 0.12), rngs)]
        steps = 6
        nccl_id = slabs.broadcast_unique_id(dist, rank, dlib.nccl_unique_id, dlib.NCCL_ID_BYTES, device="cuda")
        z0, z1 = slabs.slab_range(gn[2], rank, world)
        with Lattice(gn, 2, **cases.SC2, global_acc=[1e-5, 0, -1e-5], arithmetic="exact", kernel="auto",
                     device=local_rank, rank=rank, nranks=world, nccl_id=nccl_id) as G:
            variant = G.kernel_variant
            for k, v in opts.items():
                G.set_option(k, v)
            for c in range(2):
                G.set_populations(c, np.ascontiguousarray(pops[c][:, :, z0:z1]))
            G.step(steps)
            mine = [G.get_populations(c) for c in range(2)]
        ref = torch.empty((2,) + gn + (19,), dtype=torch.float64, device="cuda")
        if rank == 0:
            with Lattice(gn, 2, **cases.SC2, global_acc=[1e-5, 0, -1e-5], arithmetic="exact", kernel="two_pass",
                         device=local_rank) as G1:
                for c in range(2):
                    G1.set_populations(c, pops[c])
                G1.step(steps)
                for c in range(2):
                    ref[c].copy_(torch.from_numpy(G1.get_populations(c)))
        dist.broadcast(ref, 0)
        ok = all(_bitwise_equal(mine[c], np.ascontiguousarray(ref[c, :, :, z0:z1].cpu().numpy())) for c in range(2))
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        out[name] = {"bitexact": bool(t.item() == 1), "kernel": variant, "lattice": list(gn), "steps": steps,
                     "reference": "two_pass on 1 rank"}
        del ref
    out["bitexact"] = all(v["bitexact"] for v in out.values())
    return out


def run_ours(args, rank, world, local_rank):
    from dlbc_b200 import cases, init as I
    from dlbc_b200 import lib as dlib
    from dlbc_b200.lattice import Lattice

    dist = None
    nccl_id = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from dlbc_b200 import slabs
        nccl_id = slabs.broadcast_unique_id(dist, rank, dlib.nccl_unique_id, dlib.NCCL_ID_BYTES, device="cuda")

    decomp = None
    if world > 1 and not args.no_checks:
        decomp = decomposition_check(args, dist, rank, world, local_rank, dlib)
        log(f"rank {rank}: decomposition check {decomp}")

    ncomp = args.components
    gn, _ = bench_lattice(args, world)
    params = dict(cases.SC2) if ncomp == 2 else dict(tau=[1.0], eqdist_form="SecondOrder", enable_shan_chen=False)
    porous = args.workload == "porous"
    if porous:
        if ncomp != 2:
            raise SystemExit("--workload porous is the two-fluid BASELINE configs[3]")
        params["global_acc"] = [0.0, 0.0, 1e-5]
    G = Lattice(gn, ncomp, **params, arithmetic=args.arith, kernel=args.kernel, device=local_rank,
                rank=rank, nranks=world, nccl_id=nccl_id)
    for name, v in (("wave_tyb", args.wave_tyb), ("wave_la", args.wave_la)):
        if v:
            G.set_option(name, v)
    gl = G.n
    sites_local = gl[0] * gl[1] * gl[2]
    sites_total = sites_local * world

    # synthetic initial state: a few distinct pinned layer blocks, reused cyclically
    lib = dlib.load()
    layer_bytes = gl[1] * gl[2] * 19 * 8
    xc = max(1, min(gl[0], (160 << 20) // layer_bytes))
    form = I.BDIST2 if ncomp == 2 else I.SECOND_ORDER
    rho0 = 0.37 if ncomp == 2 else 1.0
    views, ptrs = [], []
    for k in range(2):
        v, p = pinned_array(lib, (xc, gl[1], gl[2], 19))
        v[...] = cases.synthetic_layers(gl, xc, rho0, 0.01, form, 42 + 7 * rank + k)
        views.append(v); ptrs.append(p)
    if porous:
        # BASELINE configs[3]: random bounce-back geometry (solid where u < 0.3, u the
        # Mt19937(1234 [+ rank]) uniform(0, 1) stream over the physical lattice in [x][y][z] order), fluids at
        # rho = 0.7 / 0.04, solid sites initialised as initEqDistWall does (fields/init.d:469-477)
        mask = cases.porous_mask(gl, 0.3, 1234 + rank)
        G.set_mask(mask)
        wall = np.zeros(19); wall[0] = 1.0
        eq = I.eq_dist_unity(cases.d3q19, form)
        for c, rho in enumerate((0.7, 0.04)):
            for x0 in range(0, gl[0], xc):
                cnt = min(xc, gl[0] - x0)
                blk = views[0][:cnt]
                blk[...] = rho * eq
                blk[mask[x0:x0 + cnt] != 0] = wall
                G.set_populations_xrange(c, blk, x0)
                G.sync()
        args.no_e2e = True  # (the e2e leg re-uploads the spinodal layers: not this workload's state)
    else:
        fill_lattice(G, gl, ncomp, views)
    G.sync()

    def barrier():
        G.sync()
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        from dlbc_b200 import slabs
        return slabs.max_over_ranks(dist, x, device="cuda")

    # ---- device-resident timing: W warm-up steps, then exactly K timed steps ----
    G.step(args.warmup)
    barrier()
    launches0 = G.launch_count
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms = G.timed_steps(args.steps)
    barrier()
    clocks = sampler.stop() if sampler else None
    launches = G.launch_count - launches0
    ms = max_over_ranks(ms)
    value = sites_total * args.steps / (ms * 1e-3) * 1e-6

    # ---- dominant kernel, CUDA events around every launch (single-rank schedule only) ----
    roof = None
    peak, peak_src = measured_peak()
    b_alg = B_ALG.get(ncomp, 2 * ncomp * 19 * 8 + 1)
    if world == 1:
        acc = {}
        reps = max(3, min(args.steps, 10))
        for _ in range(reps):
            for name, t in G.kernel_times():
                acc.setdefault(name, []).append(t)
        dom = max(acc, key=lambda k: sum(acc[k]))
        dom_ms = sum(acc[dom]) / len(acc[dom])
        achieved = b_alg * sites_local / (dom_ms * 1e-3) * 1e-9
        roof = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                "alg_bytes_per_site": b_alg, "kernel_ms": dom_ms,
                "kernels_ms": {k: sum(v) / len(v) for k, v in acc.items()},
                "step_frac": b_alg * sites_local / (ms / args.steps * 1e-3) * 1e-9 / peak,
                "frac_of_8TBs_nominal": achieved / 8000.0}
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                with open(tp) as f:
                    table = json.load(f)
                # ncu-measured DRAM bytes per site of THIS kernel instantiation (component count / mask) x the sites
                # of one launch; captured once per round on a 512x512x128 (256^3 for C = 1) launch, not in this run
                key = dom + ("/porous" if porous else "/c1" if ncomp == 1 else "")
                tr = table.get(key)
                if tr and ncomp <= 2:
                    roof["traffic"] = tr["bytes_per_site"] * sites_local
                    roof["traffic_bytes_per_site"] = tr["bytes_per_site"]
                    roof["traffic_source"] = tr["source"]
            except Exception:
                pass
        if len(acc) > 1:
            roof["note"] = ("frac is the dominant kernel's alone; the step also runs " +
                            ", ".join(k for k in acc if k != dom) + ": see step_frac")
    else:
        achieved = b_alg * sites_local / (ms / args.steps * 1e-3) * 1e-9
        roof = {"bound": "hbm", "kernel": "whole step (per GPU)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "peak_source": peak_src, "alg_bytes_per_site": b_alg}

    # ---- end to end through the C ABI with host buffers ----
    e2e = None
    if not args.no_e2e:
        import ctypes as C
        dens, dptr = pinned_array(lib, gl)
        barrier()
        t0 = time.perf_counter()
        h2d = fill_lattice(G, gl, ncomp, views)
        t_up = time.perf_counter() - t0
        G.step(args.steps, sync=False)
        d2h = 0
        for c in range(ncomp):
            G._ck(lib.dlbc_cuda_download_density(G._h, c, dens.ctypes.data_as(C.POINTER(C.c_double)), 0))
            d2h += dens.nbytes
        G.sync()
        dt = time.perf_counter() - t0
        dt = max_over_ranks(dt)
        e2e = {"value": sites_total * args.steps / dt * 1e-6, "unit": "MLUPS",
               "h2d_bytes_per_step": h2d * world / args.steps, "d2h_bytes_per_step": d2h * world / args.steps,
               "seconds": dt, "upload_seconds": t_up, "what": "pinned-host upload of all populations + K steps + download of all density fields"}
        lib.dlbc_cuda_host_free(dptr)

    # ---- the same start without host arrays: populations generated on the device ----
    # NOT the e2e number (its inputs never cross PCIe); reported beside it.
    dev_init = None
    if not args.no_e2e and not porous:
        barrier()
        t0 = time.perf_counter()
        for c in range(ncomp):
            G.init_populations(c, "EqDistPerturb", rho0, perturb=0.01, seed=42 + c + 7 * rank)
        G.sync()
        t_init = max_over_ranks(time.perf_counter() - t0)
        G.step(args.steps, sync=False)
        G.sync()
        dt = max_over_ranks(time.perf_counter() - t0)
        dev_init = {"init_seconds": t_init, "value": sites_total * args.steps / dt * 1e-6, "unit": "MLUPS", "seconds": dt,
                    "what": "dlbc_cuda_init_populations (EqDistPerturb, Mt19937 in the reference's draw order, per rank) "
                            "+ K steps; no host buffers, so not an e2e number"}

    mass = [G.total_mass(c) for c in range(ncomp)]
    variant = G.kernel_variant
    G.close()
    for p in ptrs:
        lib.dlbc_cuda_host_free(p)

    parity = None
    if world == 1 and not args.no_checks and not porous:
        parity = parity_check(args, local_rank)
        log(f"parity check {parity}")

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        m, dt, cores = cpu_port_mlups(args.cpu_sample, args.cpu_steps, 1)
        cpu = {"value": m, "unit": "MLUPS", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_sample}^3 sample of the workload, {args.cpu_steps} timesteps, {dt:.1f} s"}
        try:  # BASELINE configs[0]: merging-droplets-3d, 64^3, 4 threads standing in for `mpirun -np 4`
            m1, dt1 = cpu_config1_mlups(20)
            cpu["config0_merging_droplets_64cubed_4_threads"] = {"value": m1, "unit": "MLUPS", "seconds": dt1, "steps": 20}
        except Exception as e:  # noqa: BLE001
            cpu["config0_merging_droplets_64cubed_4_threads"] = {"error": str(e)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "MLUPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world, gn, gl, ncomp, porous),
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "device_init": dev_init, "total_mass": mass, "kernel_variant": variant, "parity_check": parity, "decomposition_invariant": decomp,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=512, help="lattice edge per GPU")
    ap.add_argument("--shape", default=None, help="nx,ny,nz per GPU (overrides --size; profiling runs)")
    ap.add_argument("--components", type=int, default=2)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default): --size^3 per GPU; strong: --size^3 in total, split into z-slabs")
    ap.add_argument("--workload", default="spinodal", choices=["spinodal", "porous"],
                    help="porous: BASELINE configs[3] (mask-heavy path), device-resident timing only")
    ap.add_argument("--arith", default="fast", choices=["fast", "exact"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "two_pass", "wave", "tma"])
    ap.add_argument("--wave-tyb", type=int, default=0, help="band height of the single-launch kernels (0: default)")
    ap.add_argument("--wave-la", type=int, default=0, help="lag of the fused items behind the psi items (0: default)")
    ap.add_argument("--cpu-sample", type=int, default=128)
    ap.add_argument("--cpu-steps", type=int, default=20)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-checks", action="store_true",
                    help="skip the parity check (N = 1) / the decomposition-invariance check (N > 1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        log(f"note: WORLD_SIZE={world} but --gpus {args.gpus}; using WORLD_SIZE")
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # `python bench.py --gpus N` without a launcher: start the N ranks ourselves (one process per GPU)
        port = 29500 + os.getpid() % 2000
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        log("re-launching under torch.distributed.run:", " ".join(cmd))
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
