"""world_size-2 tests of the multi-rank host logic on CPU (torch.distributed, gloo backend).

No GPU and no libdlbc_cuda compute calls here: these tests cover (a) the slab bookkeeping and the
torch.distributed plumbing around the C ABI (dlbc_b200/slabs.py), (b) the library's reduced ghost-plane
protocol -- only the 5 populations with c_z = +-1 cross a slab face -- on a numpy pull-streaming step, and
(c) decomposition invariance of the timestep itself: the CPU oracle run as two z-slab ranks that exchange the
reference's two halo layers over gloo (fields/parallel.d:58-103) must reproduce the one-rank run bit for bit.
The GPU versions of (c) are in tests/test_multi_gpu.py.
"""
import os
import socket

import numpy as np
import pytest

from helpers import SC, assert_bitexact


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(fn, world, *args):
    import torch.multiprocessing as mp

    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q, args)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    out = {}
    while not q.empty():
        r, ok, val = q.get()
        out[r] = (ok, val)
    for p in procs:
        assert p.exitcode == 0, f"rank process exited with {p.exitcode}: {out}"
    for r in range(world):
        assert r in out and out[r][0], f"rank {r}: {out.get(r)}"
    return [out[r][1] for r in range(world)]


def _entry(fn, rank, world, port, q, args):
    import sys
    import traceback

    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      OMP_NUM_THREADS="2")
    import torch.distributed as dist

    try:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        val = fn(dist, rank, world, *args)
        dist.barrier()
        q.put((rank, True, val))
    except Exception:
        q.put((rank, False, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


# ---- (a) plumbing -----------------------------------------------------------------------------------
def _plumbing(dist, rank, world):
    from dlbc_b200 import slabs

    lo, hi = slabs.slab_range(48, rank, world)
    assert (lo, hi) == (rank * 24, rank * 24 + 24)
    assert slabs.neighbours(rank, world) == ((rank - 1) % world, (rank + 1) % world)
    uid = slabs.broadcast_unique_id(dist, rank, lambda: bytes(range(128)), 128)
    assert uid == bytes(range(128))
    assert slabs.max_over_ranks(dist, 10.0 + rank) == 10.0 + world - 1
    return (lo, hi)


def test_slab_bookkeeping_and_id_broadcast():
    from dlbc_b200 import slabs

    assert slabs.UP_Q == (5, 9, 10, 11, 17) and slabs.DOWN_Q == (6, 12, 15, 16, 18)
    with pytest.raises(ValueError):
        slabs.slab_range(50, 0, 4)
    ranges = _run(_plumbing, 2)
    assert ranges == [(0, 24), (24, 48)]


# ---- (b) reduced ghost-plane protocol on a numpy pull step ------------------------------------------
def _pull(f_with_z_ghosts):
    """lb/advection.d:61-89 without walls: new[p][q] = old[p - c_q][q]; x, y periodic, z from ghosts."""
    from dlbc_b200.connectivity import d3q19

    a = f_with_z_ghosts
    out = np.empty_like(a[:, :, 1:-1])
    for q, (cx, cy, cz) in enumerate(d3q19.velocities):
        src = np.roll(a[..., q], (cx, cy), axis=(0, 1))
        out[..., q] = src[:, :, 1 - cz:a.shape[2] - 1 - cz]
    return out


def _reduced_protocol(dist, rank, world, gn):
    import torch
    from dlbc_b200 import slabs

    rng = np.random.default_rng(7)
    f = rng.random(gn + (19,))
    expected = _pull(np.concatenate([f[:, :, -1:], f, f[:, :, :1]], axis=2))
    lo, hi = slabs.slab_range(gn[2], rank, world)
    loc = np.full((gn[0], gn[1], hi - lo + 2, 19), np.nan)
    loc[:, :, 1:-1] = f[:, :, lo:hi]
    up, down = list(slabs.UP_Q), list(slabs.DOWN_Q)
    from_down, from_up = slabs.ring_exchange(dist, rank, world, torch.from_numpy(loc[:, :, -2, up].copy()),
                                             torch.from_numpy(loc[:, :, 1, down].copy()))
    loc[:, :, 0, up] = from_down.numpy()
    loc[:, :, -1, down] = from_up.numpy()
    got = _pull(loc)
    assert not np.isnan(got).any(), "a population outside the 5 crossing directions was read from a ghost plane"
    assert_bitexact(got, expected[:, :, lo:hi], "pull with 5-plane ghost exchange")
    return True


def test_five_crossing_populations_suffice():
    assert _run(_reduced_protocol, 2, (6, 5, 8)) == [True, True]


# ---- (c) the timestep itself: oracle as two z-slab ranks with the reference's 2-layer halo exchange ---
def _oracle_two_ranks(dist, rank, world, gn, steps):
    import torch
    from dlbc_b200 import init as I, slabs
    from dlbc_b200.connectivity import d3q19
    from oracle import oracle as O

    rng = np.random.default_rng(3)
    pops = [(rho + 0.02 * rng.random(gn))[..., None] * I.eq_dist_unity(d3q19, I.BDIST2) for rho in (0.7, 0.1)]
    lo, hi = slabs.slab_range(gn[2], rank, world)
    L = O.OracleLattice((gn[0], gn[1], hi - lo), 2, **SC, nthreads=2)
    for c, a in enumerate(pops):
        L.set_populations(c, np.ascontiguousarray(a[:, :, lo:hi]))
    for _ in range(steps):
        # L.exchangeHalo(): x then y locally (periodic), z with the ring neighbours, 2 layers, all 19
        # populations, edges and corners included (fields/parallel.d:58-103)
        for c in range(2):
            a = L.get_populations(c, halo=2)
            a = np.pad(a[2:-2, 2:-2], ((2, 2), (2, 2), (0, 0), (0, 0)), mode="wrap")
            from_down, from_up = slabs.ring_exchange(dist, rank, world, torch.from_numpy(a[:, :, -4:-2].copy()),
                                                     torch.from_numpy(a[:, :, 2:4].copy()))
            a[:, :, 0:2] = from_down.numpy()
            a[:, :, -2:] = from_up.numpy()
            L.set_populations(c, np.ascontiguousarray(a), halo=2)
        L.advect_fields(); L.reset_force(); L.add_shan_chen_force(); L.distribute_force()
        L.prepare_to_collide(); L.collide_fields()
    out = [L.get_populations(c) for c in range(2)]
    # every rank checks its own slab against a one-rank run of the whole lattice
    R = O.OracleLattice(gn, 2, **SC, nthreads=2)
    for c, a in enumerate(pops):
        R.set_populations(c, a)
    R.exchange_halo()
    R.step(steps)
    for c in range(2):
        assert_bitexact(out[c], np.ascontiguousarray(R.get_populations(c)[:, :, lo:hi]), f"rank {rank} slab, component {c}")
    return slabs.max_over_ranks(dist, float(L.total_mass(0)))


def test_oracle_decomposition_invariance_two_ranks():
    """Mirrors the reference's parallel.nc sweeps (tests/runnable: np 1 vs np 4 must agree)."""
    _run(_oracle_two_ranks, 2, (8, 6, 12), 3)
