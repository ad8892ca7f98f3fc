"""Run one of the reference's 3-D parameter files through the GPU path -- SURVEY.md 8(f) row 1.

    python -m dlbc_b200.replay case.in [--parameter name=value ...] [--out DIR]

Plays the role of src/main.d for the lattice-Boltzmann subset: parse the parameter file
(`dlbc_b200.params`), build mask and initial populations exactly as initMask / initFluids do for a
one-rank run (lb/mask.d:134-172, lb/init.d:77-136; `dlbc_b200.init`), then step and collect the fields
that dumpLBData would write (lb/io.d:56-101) at the frequencies the file asks for.  Outputs are keyed by
the reference's file-name stem, ``<field>-<simulationName>-t%08d`` (io/io.d:157-172 without the
time-stamp id), so they can be compared one to one with ``reference-data/*.h5``; ``--out`` writes them
as HDF5 files with one contiguous ``/OutArray`` dataset each (`dlbc_b200.h5lite.write_dataset`, 8(f) row 2).

Unsupported parameters fail loudly: electrokinetics, File inits, non-3-D lattices.  The timestep always
runs through libdlbc_cuda (no CPU fallback); `make_lattice` is a seam for the tests, which drive the same
driver code with their CPU oracle to check the reader and the driver without a GPU.
"""
from __future__ import annotations

import argparse
import os

import numpy as np

from . import init as I
from . import params as P
from .connectivity import d2q9, d3q19
from .rng import Mt19937

AXIS = {"X": 0, "Y": 1, "Z": 2}
CONN = {2: d2q9, 3: d3q19}


class ReplayError(ValueError):
    pass


def _fit(v, n, name, strict=False, fill=0):
    """checkArrayParameterLength (parameters.d:262-283): empty -> zeros, wrong length -> fatal."""
    v = list(v)
    if not v and n:
        if strict:
            raise ReplayError(f"Array parameter {name} must have length {n}.")
        return [fill] * n
    if len(v) != n:
        raise ReplayError(f"Array parameter {name} must have length {n}.")
    return v


def lattice_kwargs(p):
    """Parameter dict -> (gn, components, kwargs of Lattice / OracleLattice)."""
    gn = tuple(int(x) for x in p["lattice.gn"])
    if len(gn) not in (2, 3):
        raise ReplayError(f"lattice.gn = {list(gn)}: libdlbc_cuda implements D3Q19 (3-D) and D2Q9 (2-D)")
    if p["elec.enableElec"]:
        raise ReplayError("elec.enableElec = true: electrokinetics is outside the GPU path")
    C = int(p["lb.components"])
    sc = bool(p["lb.force.enableShanChen"])
    # lb/init.d:99-105: a run that never collides does not need lb.tau (the reference zero-fills it; the lattice
    # constructors want a positive value, which is then never used)
    kw = dict(tau=_fit(p["lb.tau"], C, "lb.lb.tau", strict=p["lb.timesteps"] > 0, fill=1.0),
              global_acc=_fit(p["lb.force.globalAcc"], len(gn), "lb.force.globalAcc", fill=0.0),
              psi_form=p["lb.force.psiForm"], eqdist_form=p["lb.eqdist.eqDistForm"], enable_shan_chen=sc)
    if sc:
        gcc = _fit(p["lb.force.gcc"], C, "lb.force.gcc", strict=True)
        kw["gcc"] = [_fit(row, C, f"lb.force.gcc[{i}]", strict=True) for i, row in enumerate(gcc)]
        kw["gwc"] = _fit(p["lb.force.gwc"], C, "lb.force.gwc", fill=0.0)
    return gn, C, kw


def initial_mask(p, blk, base_dir="."):
    """initMask, lb/mask.d:134-172 (one rank)."""
    kind = p["lb.mask.maskInit"]
    axis = AXIS[p["lb.mask.initAxis"]]
    if kind == "None":
        return I.mask_none(blk)
    if kind == "Box":
        return I.mask_box(blk)
    if kind == "Tube":
        return I.mask_tube(blk, axis)
    if kind == "Plates":
        return I.mask_plates(blk, axis, p["lb.mask.wallOffset"])
    if kind == "Walls":
        dhw = _fit(p["lb.mask.dimensionHasWall"], len(p["lattice.gn"]), "lb.mask.dimensionHasWall", fill=[False, False])
        return I.mask_walls(blk, [_fit(x, 2, "lb.mask.dimensionHasWall[i]", fill=False) for x in dhw],
                            p["lb.mask.wallThickness"])
    if kind == "File":  # mask.d:152-154: the global mask from an HDF5 file, relative to the run directory
        from . import h5lite
        m = I.mask_none(blk)
        m[blk.interior()] = h5lite.read_dataset(os.path.join(base_dir, p["lb.mask.maskFile"])).astype(m.dtype)
        return m
    raise ReplayError(f"lb.mask.maskInit = {kind} is not supported by the replay driver")


def rank_grid(p, gn):
    """The process grid the parameter file asks for (parallel.nc, parallel.d:190-237) as [(rank, M.c)], ranks in the
    row-major order of MPI_Cart_create; a grid with a 0 (MPI_Dims_create decides) or a missing entry is one rank."""
    nc = [int(v) for v in p.get("parallel.nc", [])]
    if len(nc) != len(gn) or any(v < 1 for v in nc) or any(g % v for g, v in zip(gn, nc)):
        nc = [1] * len(gn)
    return tuple(nc), [(r, c) for r, c in enumerate(np.ndindex(*nc))]


def _per_rank(p, blk, make):
    """A RANDOM initialiser draws one number per site of every rank's own array (halo included), each rank from its
    own generator (lb/init.d:109-118, random.d:43-53): emulate the ranks of parallel.nc one after the other and put
    their blocks together.  make(rank block, rank) -> that rank's array."""
    nc, ranks = rank_grid(p, blk.gn)
    if len(ranks) == 1:
        return make(blk, 0)
    out = None
    for r, c in ranks:
        b = I.Block(blk.gn, halo=blk.halo, nc=nc, c=c)
        a = make(b, r)
        if out is None:
            out = np.zeros(blk.shape_h + a.shape[len(blk.gn):], dtype=a.dtype)
        dst = tuple(slice(blk.halo + ci * n, blk.halo + (ci + 1) * n) for ci, n in zip(c, b.n))
        out[dst] = a[b.interior()]
    return out


def initial_populations(p, blk, mask, base_dir="."):
    """initFluids, lb/init.d:77-136.  Deterministic initialisers are functions of the global coordinates; the random
    ones are generated rank by rank for the process grid of the parameter file (`_per_rank`)."""
    C = int(p["lb.components"])
    form = I.EQDIST_FORMS[p["lb.eqdist.eqDistForm"]]
    kinds = _fit(p["lb.init.fluidInit"], C, "lb.init.fluidInit", fill="None")
    dens = _fit(p["lb.init.fluidDensities"], C, "lb.init.fluidDensities", fill=[])
    perturb = _fit(p["lb.init.fluidPerturb"], C, "lb.init.fluidPerturb", fill=0.0)
    reseed = _fit(p["lb.init.randomReseed"], C, "lb.init.randomReseed", fill=False)
    seeds = _fit(p["lb.init.randomSeed"], C, "lb.init.randomSeed", fill=0)
    nd = len(p["lattice.gn"])
    conn = CONN[nd]  # the connectivity of this lattice (gconn, connectivity.d:21-32): conn, or d2q9 on a 2-D lattice
    off = _fit(p["lb.init.initOffset"], nd, "lb.init.initOffset", fill=0.0)
    sep = _fit(p["lb.init.initSeparation"], nd, "lb.init.initSeparation", fill=0.0)
    axis = AXIS[p["lb.init.initAxis"]]
    radius, thick = p["lb.init.initRadius"], p["lb.init.interfaceThickness"]
    widths = list(p["lb.init.lamellaeWidths"])
    shift = bool(p.get("random.shiftSeedByRank", False))
    nranks = len(rank_grid(p, blk.gn)[1])
    # the global generator of every rank (random.d:43-53), shared by that rank's non-reseeding fluids
    rngs = [Mt19937(p["random.seed"] + (r if shift else 0)) for r in range(nranks)]
    pops = []
    for f in range(C):
        k, d = kinds[f], list(dens[f])

        def random_init(fn):
            def make(b, r):
                lrng = Mt19937(seeds[f] + (r if shift else 0)) if reseed[f] else rngs[r]
                return fn(b, lrng)
            return _per_rank(p, blk, make)

        def need(n):
            if len(d) != n:
                raise ReplayError(f"Array parameter lb.init.fluidDensities[{f}] must have length {n}.")
        if k == "None":  # lb/init.d:270-271: the field stays as allocated -- double.init, a quiet NaN
            a = np.full(blk.shape_h + (conn.q,), np.nan)
        elif k == "File":  # :272-275: the global populations from an HDF5 file, relative to the run directory
            from . import h5lite
            files = _fit(p["lb.init.fluidFiles"], C, "lb.init.fluidFiles", strict=True)
            a = np.zeros(blk.shape_h + (conn.q,))
            a[blk.interior()] = h5lite.read_dataset(os.path.join(base_dir, files[f]))
        elif k == "Const":
            need(1); a = I.init_const(blk, conn.q, d[0])
        elif k == "ConstRandom":
            need(1); a = random_init(lambda b, g: I.init_const_random(b, conn.q, d[0], g))
        elif k == "EqDist":
            need(1); a = I.init_eq_dist(blk, conn, form, d[0])
        elif k == "EqDistRandom":
            need(1); a = random_init(lambda b, g: I.init_eq_dist_random(b, conn, form, d[0], g))
        elif k == "EqDistPerturb":
            need(1); a = random_init(lambda b, g: I.init_eq_dist_perturb(b, conn, form, d[0], perturb[f], g))
        elif k == "EqDistPerturbFrac":
            need(1); a = random_init(lambda b, g: I.init_eq_dist_perturb_frac(b, conn, form, d[0], perturb[f], g))
        elif k == "EqDistSphere":
            need(2); a = I.init_eq_dist_sphere(blk, conn, form, d[0], d[1], radius, off, thick)
        elif k == "EqDistSphereFrac":
            need(2); a = I.init_eq_dist_sphere_frac(blk, conn, form, d[0], d[1], radius, off, thick)
        elif k == "EqDistTwoSpheres":
            need(2); a = I.init_eq_dist_two_spheres(blk, conn, form, d[0], d[1], radius, off, thick, sep)
        elif k == "EqDistTwoSpheresFrac":
            need(2); a = I.init_eq_dist_two_spheres_frac(blk, conn, form, d[0], d[1], radius, off, thick, sep)
        elif k == "EqDistCylinder":
            need(2); a = I.init_eq_dist_cylinder(blk, conn, form, d[0], d[1], axis, radius, off, thick)
        elif k == "EqDistCylinderFrac":
            need(2); a = I.init_eq_dist_cylinder_frac(blk, conn, form, d[0], d[1], axis, radius, off, thick)
        elif k == "EqDistLamellae":
            need(len(widths)); a = I.init_eq_dist_lamellae(blk, conn, form, d, widths, axis, thick)
        elif k == "EqDistLamellaeFrac":
            need(len(widths)); a = I.init_eq_dist_lamellae_frac(blk, conn, form, d, widths, axis, thick)
        else:
            raise ReplayError(f"lb.init.fluidInit[{f}] = {k} is not supported by the replay driver")
        pops.append(I.init_eq_dist_wall(np.ascontiguousarray(a), mask))  # lb/init.d:121
    return pops


def device_initialise(L, p):
    """initMask + initFluids + initEqDistWall generated ON THE DEVICE (dlbc_cuda_init_*): no host
    array of the lattice's size ever exists.  One-rank semantics as `initial_populations`; raises ReplayError for the
    initialisers the device does not generate (Const, ConstRandom, File)."""
    C = int(p["lb.components"])
    gn = tuple(int(x) for x in p["lattice.gn"])
    kind = p["lb.mask.maskInit"]
    if kind not in ("None", "Box", "Tube", "Plates", "Walls"):
        raise ReplayError(f"lb.mask.maskInit = {kind} is not generated on the device")
    dhw = None
    if kind == "Walls":
        dhw = [_fit(x, 2, "lb.mask.dimensionHasWall[i]", fill=False)
               for x in _fit(p["lb.mask.dimensionHasWall"], 3, "lb.mask.dimensionHasWall", fill=[False, False])]
    L.init_mask(kind, axis=AXIS[p["lb.mask.initAxis"]], wall_offset=p["lb.mask.wallOffset"],
                thickness=p["lb.mask.wallThickness"], dimension_has_wall=dhw)
    kinds = _fit(p["lb.init.fluidInit"], C, "lb.init.fluidInit", fill="None")
    dens = _fit(p["lb.init.fluidDensities"], C, "lb.init.fluidDensities", fill=[])
    perturb = _fit(p["lb.init.fluidPerturb"], C, "lb.init.fluidPerturb", fill=0.0)
    reseed = _fit(p["lb.init.randomReseed"], C, "lb.init.randomReseed", fill=False)
    seeds = _fit(p["lb.init.randomSeed"], C, "lb.init.randomSeed", fill=0)
    off = _fit(p["lb.init.initOffset"], 3, "lb.init.initOffset", fill=0.0)
    sep = _fit(p["lb.init.initSeparation"], 3, "lb.init.initSeparation", fill=0.0)
    axis = AXIS[p["lb.init.initAxis"]]
    radius, thick = p["lb.init.initRadius"], p["lb.init.interfaceThickness"]
    widths = list(p["lb.init.lamellaeWidths"])
    nfull = (gn[0] + 4) * (gn[1] + 4) * (gn[2] + 4)  # draws one random initialiser takes from its generator
    shared_draws = 0                                   # consumed so far from the global generator (random.d:43-53)
    for f in range(C):
        k, d = kinds[f], list(dens[f])
        frac = k.endswith("Frac")
        if k in ("EqDist", "EqDistRandom", "EqDistPerturb", "EqDistPerturbFrac"):
            seed, skip = (seeds[f], 0) if reseed[f] else (p["random.seed"], shared_draws)
            L.init_populations(f, k, d[0], perturb=perturb[f], seed=seed, skip=skip)
            if k != "EqDist" and not reseed[f]:
                shared_draws += nfull
        elif k in ("EqDistSphere", "EqDistSphereFrac"):
            L.init_spheres(f, d[0], d[1], radius, thick, offset=off, frac=frac)
        elif k in ("EqDistTwoSpheres", "EqDistTwoSpheresFrac"):
            L.init_spheres(f, d[0], d[1], radius, thick, offset=off, separation=sep, frac=frac)
        elif k in ("EqDistCylinder", "EqDistCylinderFrac"):
            L.init_spheres(f, d[0], d[1], radius, thick, offset=off, cylinder_axis=axis, frac=frac)
        elif k in ("EqDistLamellae", "EqDistLamellaeFrac"):
            L.init_lamellae(f, d, widths, axis, thick, frac=frac)
        else:
            raise ReplayError(f"lb.init.fluidInit[{f}] = {k} is not generated on the device")
    L.init_wall()  # lb/init.d:121


def _dump_now(freq, t):
    """io/io.d:224-226"""
    return freq > 0 and t % freq == 0


def _collect(L, p, t, out):
    """dumpLBData, lb/io.d:56-101 (laplace output excluded)."""
    if not p["io.enableIO"] or t < p["io.startOutput"]:  # io/io.d dumpData
        return
    names = list(p["lb.fieldNames"])
    C = int(p["lb.components"])
    sim = P.replace_fname_tokens(p["io.simulationName"], p)  # e.g. laplace-droplet-3d-gcc%lb.force.gcc[1][0]%

    def put(stem, arr):
        out[f"{stem}-{sim}-t{t:08d}"] = arr

    io = lambda k: p[f"lb.io.{k}Freq"]
    if _dump_now(io("population"), t):
        for i in range(C):
            put(f"population-{names[i]}", L.get_populations(i))
    if _dump_now(io("density"), t):
        for i in range(C):
            put(f"density-{names[i]}", L.density(i))
    if _dump_now(io("colour"), t):
        for i in range(C):
            for j in range(i + 1, C):
                put(f"colour-{names[i]}-{names[j]}", L.colour(i, j))
    if _dump_now(io("velocity"), t):
        for i in range(C):
            put(f"velocity-{names[i]}", L.velocity(i))
    if _dump_now(io("force"), t):
        for i in range(C):
            put(f"force-{names[i]}", L.force(i))
    if _dump_now(io("mask"), t):
        put("mask", L.get_mask())


def checkpoint_paths(p, restore, base_dir="."):
    """readCheckpoint's file names (io/checkpoint.d:80-103,217-241): <outputPath>/<cpRestorePath>/cp-fluids<i>-<name>.h5
    and cp-mask-<name>.h5, relative to the directory the reference would be started in."""
    C = int(p["lb.components"])
    d = os.path.normpath(os.path.join(base_dir, p["io.outputPath"], p["io.checkpoint.cpRestorePath"]))
    return [os.path.join(d, f"cp-fluids{i}-{restore}.h5") for i in range(C)], os.path.join(d, f"cp-mask-{restore}.h5")


def read_checkpoint(p, restore, base_dir="."):
    """-> (mask int32 [x][y][z], populations per fluid [x][y][z][19], timestep) of checkpoint `restore`
    (the reference's `-r <name>`; dlbc.lb.lb.timestep comes from the `globals` group, checkpoint.d:248-313)."""
    from . import h5lite
    fluids, maskf = checkpoint_paths(p, restore, base_dir)
    pops = [h5lite.read_dataset(f) for f in fluids]
    mask = h5lite.read_dataset(maskf).astype(np.int32)
    t0 = int(h5lite.read_attributes(fluids[0], "globals")["dlbc.lb.lb.timestep"])
    return mask, pops, t0


def write_checkpoint(L, p, t, directory, input_lines=None, run_id="00000000T000000"):
    """dumpCheckpoint (io/checkpoint.d:191-215): cp-fluids<i>-<sim>-<id>-t%08d.h5 for every fluid and cp-mask-... for
    the mask under <directory>/<cpPath>, each with the `input` attribute, `metadata` and `globals/dlbc.lb.lb.timestep`,
    so that the reference's `-r <sim>-<id>-t%08d` (or this module's) restarts from them."""
    from . import h5lite
    sim = P.replace_fname_tokens(p["io.simulationName"], p)
    d = os.path.join(directory, p["io.checkpoint.cpPath"])
    os.makedirs(d, exist_ok=True)
    meta = metadata()
    paths = []
    for i in range(int(p["lb.components"])):
        f = L.get_populations(i)
        paths.append(h5lite.write_file(os.path.join(d, f"cp-fluids{i}-{sim}-{run_id}-t{t:08d}.h5"), f, input_lines=input_lines,
                                       metadata=meta, globals_={"dlbc.lb.lb.timestep": int(t)}, chunks=chunk_block(p, f.shape)))
    m = L.get_mask().astype(np.int32)
    paths.append(h5lite.write_file(os.path.join(d, f"cp-mask-{sim}-{run_id}-t{t:08d}.h5"), m, input_lines=input_lines,
                                   metadata=meta, globals_={"dlbc.lb.lb.timestep": int(t)}, chunks=chunk_block(p, m.shape)))
    return paths


def metadata():
    """The `metadata` group (io/hdf5.d:290-299): revision of the code that wrote the file."""
    return {"revisionHash": "dlbc_b200", "revisionDesc": "libdlbc_cuda (DLBC timestep on B200)", "revisionBranch": "main",
            "revisionChanged": 0, "revisionChanges": ""}


def _cuda_lattice(gn, C, kw, device=0, arithmetic="fast", kernel="auto"):
    from .lattice import Lattice
    return Lattice(gn, C, **kw, device=device, arithmetic=arithmetic, kernel=kernel)


def replay(p, make_lattice=None, device_init=False, restore=None, base_dir=".", checkpoint_dir=None, input_lines=None,
           **lattice_options):
    """Run the case described by parameter dict `p`; returns {file-name stem: array}.
    make_lattice(gn, components, kwargs) -> object with the `Lattice` interface (default: the CUDA library).
    device_init: generate mask and initial populations on the device instead of uploading host arrays.
    restore: name of a checkpoint to start from (the reference's `-r`), looked up relative to `base_dir`.
    checkpoint_dir: write checkpoints there every io.checkpoint.cpFreq steps (dumpCheckpoint)."""
    gn, C, kw = lattice_kwargs(p)
    names = _fit(p["lb.fieldNames"], C, "lb.lb.fieldNames", strict=True)
    p = dict(p, **{"lb.fieldNames": names})
    blk = I.Block(gn)
    t_start = 0
    if restore is not None:
        mask0, pops0, t_start = read_checkpoint(p, restore, base_dir)
        blk = I.Block(gn, halo=0)
        mask, pops, device_init = mask0, pops0, False
    elif not device_init:
        mask = initial_mask(p, blk, base_dir)
        pops = initial_populations(p, blk, mask, base_dir)
    L = make_lattice(gn, C, kw) if make_lattice else _cuda_lattice(gn, C, kw, **lattice_options)
    out = {}
    try:
        if device_init:
            device_initialise(L, p)
        else:
            L.set_mask(mask, halo=blk.halo)
            for c, a in enumerate(pops):
                L.set_populations(c, a, halo=blk.halo)
        if hasattr(L, "exchange_halo"):
            L.exchange_halo()  # main.d:142 (the CUDA library maintains its ghost planes itself)
        _collect(L, p, t_start, out)  # "First data dump", main.d:149 (after a restore: at the restored timestep)
        freqs = [p[f"lb.io.{k}Freq"] for k in ("population", "density", "colour", "velocity", "force", "mask")]
        cp_freq = int(p["io.checkpoint.cpFreq"]) if checkpoint_dir else 0
        freqs = sorted({f for f in freqs + [cp_freq] if f > 0})
        t, T = t_start, int(p["lb.timesteps"])
        while t < T:
            nxt = min([T] + [(t // f + 1) * f for f in freqs])  # next step at which anything is dumped
            L.step(nxt - t)
            t = nxt
            _collect(L, p, t, out)
            if cp_freq and t % cp_freq == 0:
                write_checkpoint(L, p, t, checkpoint_dir, input_lines)
    finally:
        if hasattr(L, "close"):
            L.close()
    return out


def chunk_block(p, shape):
    """The chunk of a `hdf5.writeChunked = true` file (io/hdf5.d:150-200,237-241): one rank's block -- the global
    extent divided by the process grid `parallel.nc` (a 0 or missing entry counts as 1 here: this library's ranks are
    z-slabs and the replay driver is one process) -- times the whole vector length of a vector field.  None when
    the parameter set asks for contiguous files."""
    if p is None or not p.get("io.hdf5.writeChunked", False):
        return None
    nc = [int(v) if int(v) > 0 else 1 for v in p.get("parallel.nc", [])]
    nc += [1] * (3 - len(nc))
    return tuple(-(-int(n) // nc[d]) if d < 3 else int(n) for d, n in enumerate(shape))


def write_outputs(out, directory, run_id="00000000T000000", input_lines=None, params=None):
    """Write every collected field as the reference would name and furnish it (io/io.d:157-172, io/hdf5.d:264-299):
    <field>-<sim>-<run id>-t%08d.h5 with one /OutArray dataset (contiguous, or chunked by rank block when `params`
    has hdf5.writeChunked = true), the `input` attribute (the parameter file, so that any output file can serve as
    input file again) and the `metadata` group."""
    from . import h5lite
    os.makedirs(directory, exist_ok=True)
    paths = []
    for k, v in out.items():
        stem, t = k.rsplit("-t", 1)
        paths.append(h5lite.write_file(os.path.join(directory, f"{stem}-{run_id}-t{t}.h5"), v, input_lines=input_lines,
                                       metadata=metadata() if input_lines is not None else None,
                                       chunks=chunk_block(params, np.shape(v))))
    return paths


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("parameter_file")
    ap.add_argument("-p", "--parameter", action="append", default=[], metavar="name=value",
                    help="override, as the reference's --parameter (repeatable)")
    ap.add_argument("--out", default=None, help="directory for the HDF5 files (/OutArray, the reference's names)")
    ap.add_argument("--run-id", default="00000000T000000", help="time-stamp id in the file names (io/io.d:157-172)")
    ap.add_argument("--arith", default="fast", choices=["fast", "exact"])
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("-r", "--restore", default=None, metavar="NAME",
                    help="restart from checkpoint NAME (<outputPath>/<cpRestorePath>/cp-fluids<i>-NAME.h5, cp-mask-NAME.h5), "
                         "as the reference's -r")
    ap.add_argument("--device-init", action="store_true",
                    help="generate mask and initial populations on the GPU (dlbc_cuda_init_*) instead of uploading them")
    a = ap.parse_args(argv)
    p = P.parse_file(a.parameter_file, a.parameter)  # (an .h5 written by DLBC or by --out works as parameter file too)
    lines = P.read_lines(a.parameter_file)
    out = replay(p, device_init=a.device_init, restore=a.restore, base_dir=os.path.dirname(os.path.abspath(a.parameter_file)),
                 checkpoint_dir=a.out, input_lines=lines, arithmetic=a.arith, device=a.device)
    for k, v in out.items():
        print(f"{k}: shape {v.shape}, sum {float(np.sum(v)):.15g}")
    if a.out:
        write_outputs(out, a.out, a.run_id, input_lines=lines, params=p)


if __name__ == "__main__":
    main()
