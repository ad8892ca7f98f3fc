"""dlbc_cuda_mass_momentum_no_alloc (include/dlbc_cuda.h, csrc/no_alloc.cu): the allocation-free, context-free
variant of the mass / momentum reduction, called through the C ABI with device pointers and checked against the
existing entry points (dlbc_cuda_total_mass / dlbc_cuda_total_momentum) on the same data.  Bit-identical by
construction (same grid, same tree order).  Collected last on purpose; passes
alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from helpers import random_case

MAXC = 4            # DLBC_CUDA_NO_ALLOC_MAX_COMPONENTS
MAX_SITES = 1 << 31  # DLBC_CUDA_NO_ALLOC_MAX_SITES


def _call(lib, ncomp, gn, d_pops, d_mask, d_out, stream=None):
    err = C.create_string_buffer(256)
    rc = lib.dlbc_cuda_mass_momentum_no_alloc(ncomp, gn[0], gn[1], gn[2], d_pops, d_mask, d_out, stream, err, len(err))
    return rc, err.value.decode()


def test_no_alloc_rejects_sizes_beyond_the_maximum():
    """Argument checks come before any CUDA call: runs without a GPU."""
    from dlbc_b200 import lib as LIB
    L = LIB.load()
    rc, msg = _call(L, MAXC + 1, (8, 8, 8), 1, None, 1)
    assert rc != 0 and "beyond the maximum" in msg and "ncomp = 5" in msg
    rc, msg = _call(L, 1, (2048, 2048, 513), 1, None, 1)
    assert rc != 0 and "beyond the maximum" in msg and "2048 x 2048 x 513" in msg
    rc, msg = _call(L, 1, (1 << 40, 1 << 40, 1 << 40), 1, None, 1)
    assert rc != 0 and "beyond the maximum" in msg
    rc, msg = _call(L, 0, (8, 8, 8), 1, None, 1)
    assert rc != 0
    rc, msg = _call(L, 1, (8, 8, 8), None, None, 1)
    assert rc != 0 and "null" in msg
    # a NULL message buffer is allowed
    assert L.dlbc_cuda_mass_momentum_no_alloc(MAXC + 1, 8, 8, 8, 1, None, 1, None, None, 0) != 0


def _existing(gn, solid, seed, steps):
    """The existing entry points on a stepped lattice: (populations [c][x][y][z][19], mask, [c][4] sums)."""
    from dlbc_b200.lattice import Lattice
    params, pops, mask = random_case(gn, solid, seed=seed)
    p = dict(params); p.pop("gn"); nc = p.pop("components")
    with Lattice(gn, nc, **p, arithmetic="exact") as G:
        if mask is not None:
            G.set_mask(mask)
        for c, a in enumerate(pops):
            G.set_populations(c, a)
        G.step(steps)
        got = [G.get_populations(c) for c in range(nc)]
        sums = np.array([[G.total_mass(c), *G.total_momentum(c)] for c in range(nc)])
    return got, mask, sums


def _device_inputs(torch, pops, mask):
    """[c][x][y][z][19] -> device [c][19][z][y][x]; int32 mask [x][y][z] -> device uint8 [z][y][x]."""
    a = np.ascontiguousarray(np.stack(pops).transpose(0, 4, 3, 2, 1))
    d_pops = torch.from_numpy(a).cuda()
    d_mask = None if mask is None else torch.from_numpy(np.ascontiguousarray((mask != 0).astype(np.uint8).transpose(2, 1, 0))).cuda()
    return d_pops, d_mask


@pytest.mark.gpu
def test_no_alloc_matches_existing_entry_points_on_calls_in_a_row():
    import torch
    from dlbc_b200 import lib as LIB
    L = LIB.load()
    # case A: two components with solid sites, ragged sizes; case B: no mask, another size (fewer sites than
    # threads in the grid, so most blocks must overwrite last call's partials with zeros)
    gn_a, gn_b = (36, 20, 14), (12, 10, 6)
    pops_a, mask_a, want_a = _existing(gn_a, 0.2, seed=5, steps=3)
    pops_b, mask_b, want_b = _existing(gn_b, 0.0, seed=8, steps=2)
    assert mask_a is not None and mask_b is None
    da, ma = _device_inputs(torch, pops_a, mask_a)
    db, mb = _device_inputs(torch, pops_b, mask_b)
    out = torch.full((3, 2, 4), float("nan"), dtype=torch.float64, device="cuda")
    side = torch.cuda.Stream()
    torch.cuda.synchronize()

    def run(i, gn, d, m, stream=None):
        rc, msg = _call(L, 2, gn, d.data_ptr(), None if m is None else m.data_ptr(), out[i].data_ptr(), stream)
        assert rc == 0, msg

    # two calls in a row (A then B) on the default stream, then A again on another stream once those are done;
    # the scratch array still holds B's partials when the third call starts
    run(0, gn_a, da, ma)
    run(1, gn_b, db, mb)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    run(2, gn_a, da, ma, side.cuda_stream)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the call changed the amount of free device memory"
    got = out.cpu().numpy()
    # bit-identical: the same grid and reduction tree as field_io.cu (the tests of total_mass use 1e-14 relative)
    for name, g, w in (("A", got[0], want_a), ("B after A", got[1], want_b), ("A after B, side stream", got[2], want_a)):
        assert np.array_equal(g, w), f"{name}: no_alloc {g.tolist()} != existing {w.tolist()}"
    # a single component of a multi-component array: the second component's sums through a pointer offset
    one = torch.zeros(4, dtype=torch.float64, device="cuda")
    sites = gn_a[0] * gn_a[1] * gn_a[2]
    rc, msg = _call(L, 1, gn_a, da.data_ptr() + 19 * sites * 8, ma.data_ptr(), one.data_ptr())
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert np.array_equal(one.cpu().numpy(), want_a[1])
