"""Decomposition invariance of the z-slab multi-GPU path: the same case on 1 GPU and
split over G GPUs (NCCL ghost-plane exchange) must give bit-identical populations, because the
per-site arithmetic does not depend on the split.  This mirrors the reference's `parallel.nc`
sweeps (tests/runnable/dlbct: every test is run on several process grids and compared with the
same golden data).  One host thread drives one context, as the C ABI prescribes."""
import threading

import numpy as np
import pytest

from helpers import assert_bitexact, assert_close, random_case

pytestmark = pytest.mark.gpu

cases = pytest.importorskip("dlbc_b200.cases")
from dlbc_b200 import lib as dlib  # noqa: E402
from dlbc_b200.lattice import Lattice, slab_range  # noqa: E402
from oracle import oracle as O  # noqa: E402


def _ngpu():
    try:
        return dlib.device_count()
    except Exception:
        return 0


def run_slabs(params, pops, mask, nranks, steps, options=None, **kw):
    """Run `steps` timesteps on `nranks` GPUs; returns the gathered global populations, masses."""
    p = dict(params)
    gn, nc = p.pop("gn"), p.pop("components")
    nccl_id = dlib.nccl_unique_id() if nranks > 1 else None
    out = [None] * nranks
    err = []

    def worker(r):
        try:
            z0, z1 = slab_range(gn[2], r, nranks)
            with Lattice(gn, nc, **p, device=r, rank=r, nranks=nranks, nccl_id=nccl_id, **kw) as G:
                for k, v in (options or {}).items():
                    G.set_option(k, v)
                if mask is not None:
                    G.set_mask(np.ascontiguousarray(mask[:, :, z0:z1]))
                for c, a in enumerate(pops):
                    G.set_populations(c, np.ascontiguousarray(a[:, :, z0:z1]))
                G.step(steps)
                res = [G.get_populations(c) for c in range(nc)]
                dens = [G.density(c) for c in range(nc)]
                mass = [G.total_mass(c) for c in range(nc)]
                out[r] = (res, dens, mass)
        except Exception as e:  # noqa: BLE001
            err.append((r, e))

    ts = [threading.Thread(target=worker, args=(r,), daemon=True) for r in range(nranks)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=90)
    assert not err, err
    assert all(o is not None for o in out), "a rank did not finish (hang?)"
    f = [np.concatenate([out[r][0][c] for r in range(nranks)], axis=2) for c in range(nc)]
    d = [np.concatenate([out[r][1][c] for r in range(nranks)], axis=2) for c in range(nc)]
    return f, d, out[0][2]


@pytest.mark.parametrize("kernel", ["two_pass", "wave", "tma"])
@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_decomposition_invariance_spinodal(kernel, nranks):
    if _ngpu() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    params, pops, _ = cases.spinodal((32, 24, 48))
    f1, d1, m1 = run_slabs(params, pops, None, 1, 6, arithmetic="exact", kernel=kernel)
    fg, dg, mg = run_slabs(params, pops, None, nranks, 6, arithmetic="exact", kernel=kernel)
    for c in range(2):
        assert_bitexact(fg[c], f1[c], f"{kernel} x{nranks}: populations {c}")
        assert_bitexact(dg[c], d1[c], f"{kernel} x{nranks}: density {c}")
        assert abs(mg[c] - m1[c]) <= 1e-13 * abs(m1[c])


@pytest.mark.parametrize("kernel", ["two_pass", "wave"])
def test_decomposition_invariance_porous_wall_term(kernel):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    params, pops, mask = cases.porous((24, 20, 16), gwc=[-0.05, 0.05])
    f1, _, _ = run_slabs(params, pops, mask, 1, 5, arithmetic="exact", kernel=kernel)
    fg, _, _ = run_slabs(params, pops, mask, 2, 5, arithmetic="exact", kernel=kernel)
    for c in range(2):
        assert_bitexact(fg[c], f1[c], f"{kernel}: porous populations {c}")


@pytest.mark.parametrize("kernel", ["wave", "tma"])
@pytest.mark.parametrize("nranks", [2, 4])
def test_decomposition_invariance_multi_band_slabs(kernel, nranks):
    """The single-launch schedules on z-slabs in the multi-band regime (4 bands of 8 rows + a ragged one,
    minimum lag, 3 x-segments for the TMA kernel), with a mask on the register path: 1 GPU, G slabs and the two-kernel schedule
    must agree bit for bit."""
    if _ngpu() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    params, pops, mask = random_case((136, 36, 8 * nranks), solid=0.0 if kernel == "tma" else 0.1, seed=21)
    opts = dict(wave_tyb=8, wave_la=2)
    f1, _, _ = run_slabs(params, pops, mask, 1, 6, arithmetic="exact", kernel="two_pass")
    fs, _, _ = run_slabs(params, pops, mask, 1, 6, options=opts, arithmetic="exact", kernel=kernel)
    fg, _, _ = run_slabs(params, pops, mask, nranks, 6, options=opts, arithmetic="exact", kernel=kernel)
    for c in range(2):
        assert_bitexact(fs[c], f1[c], f"{kernel} 1 GPU vs two_pass: populations {c}")
        assert_bitexact(fg[c], f1[c], f"{kernel} x{nranks} vs two_pass on 1 GPU: populations {c}")


def test_two_slabs_match_oracle():
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    params, pops, _ = cases.spinodal((16, 16, 32))
    p = dict(params)
    gn, nc = p.pop("gn"), p.pop("components")
    R = O.OracleLattice(gn, nc, **p)
    for c, a in enumerate(pops):
        R.set_populations(c, a)
    R.exchange_halo()
    R.step(8)
    fg, dg, _ = run_slabs(params, pops, None, 2, 8)
    for c in range(nc):
        assert_close(fg[c], R.get_populations(c), what=f"2 slabs vs oracle: populations {c}")
        assert_close(dg[c], R.density(c), what=f"2 slabs vs oracle: density {c}")
