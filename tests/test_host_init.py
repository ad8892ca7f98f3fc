"""Host-side setup (RNG, initial conditions, masks, HDF5 reader) against the reference's own
known answers: random.d:56-87 KATs, lb/init-2d goldens, t=0 goldens of the 3-D tests."""
import numpy as np
import pytest

from dlbc_b200 import init as I
from dlbc_b200.connectivity import d2q9, d3q19
from dlbc_b200.rng import Mt19937
from helpers import assert_bitexact


def test_mt19937_uniform_known_answers():
    # src/dlbc/random.d:56-87
    kat = {
        0: [1.0976270048640728305144875776022672653198242187500000,
            1.1856892330538686408658577420283108949661254882812500,
            1.4303787302762218658358506218064576387405395507812500],
        1: [0.8340439970684339066053780697984620928764343261718750,
            1.9943696167306901312343825338757596909999847412109375,
            1.4406489789114911292955412136507220566272735595703125],
        42: [0.7490802287936861869610538633423857390880584716796875,
             1.5930859687722020989752991226851008832454681396484375,
             1.9014286235676676195538448155275546014308929443359375],
    }
    for seed, vals in kat.items():
        r = Mt19937(seed)
        assert [float(r.uniform(0.0, 2.0)) for _ in range(3)] == vals
        r.seed(seed)  # fresh seeding starts from the top again
        assert list(r.uniform(0.0, 2.0, 3)) == vals


def test_connectivity_tables():
    # lb/connectivity.d:432-480
    for conn in (d3q19, d2q9):
        for q in range(conn.q):
            assert (conn.velocities[q] == -conn.velocities[conn.bounce[q]]).all()
        assert abs(conn.weights.sum() - 1.0) < 1e-15
        assert conn.css == 1.0 / 3.0
    assert list(d3q19.bounce) == [0, 2, 1, 4, 3, 6, 5, 14, 13, 16, 15, 18, 17, 8, 7, 10, 9, 12, 11]


def _assemble_2x2(make):
    """lb/init-2d runs with np = 4, nc = [2, 2]; rank r has M.c = (r // 2, r % 2)."""
    out = None
    for cx in range(2):
        for cy in range(2):
            blk = I.Block((64, 64), nc=(2, 2), c=(cx, cy))
            a = make(blk)[blk.interior()]
            if out is None:
                out = np.empty((64, 64) + a.shape[2:])
            out[cx * 32:(cx + 1) * 32, cy * 32:(cy + 1) * 32] = a
    return out


@pytest.mark.parametrize("name", ["eqdistperturb", "eqdistperturbfrac", "eqdistrandom", "constrandom"])
def test_random_inits_match_reference(golden, name):
    g = golden("init-2d-random")
    for comp, (colour, rho, seed) in enumerate((("red", 0.5, 42), ("blue", 1.0, 43))):
        def make(blk):
            lrng = Mt19937(seed)  # shiftSeedByRank = false, randomReseed = true (lb/init.d:109-118)
            if name == "eqdistperturb":
                return I.init_eq_dist_perturb(blk, d2q9, I.BDIST2, rho, (0.1, 0.2)[comp], lrng)
            if name == "eqdistperturbfrac":
                return I.init_eq_dist_perturb_frac(blk, d2q9, I.BDIST2, rho, 0.2, lrng)
            if name == "eqdistrandom":
                return I.init_eq_dist_random(blk, d2q9, I.BDIST2, rho, lrng)
            return I.init_const_random(blk, 9, rho, lrng)
        assert_bitexact(_assemble_2x2(make), g[f"{name}-{colour}"], f"{name} {colour}")


def test_mask_plates_matches_poiseuille_golden(golden):
    g = golden("poiseuille-2d")
    blk = I.Block((32, 4))
    m = I.mask_plates(blk, 0, 0)
    assert_bitexact(m[blk.interior()], g["mask-t0"], "mask")


def test_masks_basic():
    blk = I.Block((6, 5, 4))
    box = I.mask_box(blk)[blk.interior()]
    assert box.sum() == 6 * 5 * 4 - 4 * 3 * 2
    tube = I.mask_tube(blk, 2)[blk.interior()]
    assert tube.sum() == (6 * 5 - 4 * 3) * 4
    walls = I.mask_walls(blk, [[True, False], [False, False], [False, True]], 2)[blk.interior()]
    assert walls[:2].all() and walls[:, :, -2:].all() and not walls[2:, :, :-2].any()
    with pytest.raises(ValueError):
        I.mask_plates(I.Block((4, 4)), 2)


def test_eq_dist_unity():
    assert_bitexact(I.eq_dist_unity(d3q19, I.BDIST2), d3q19.weights, "bdist2 unity == weights (D3Q19)")
    assert_bitexact(I.eq_dist_unity(d3q19, I.SECOND_ORDER), d3q19.weights)
    u = I.eq_dist_unity(d2q9, I.BDIST2)
    assert u[0] == 0.5 and u[1] == 2.0 * (1.0 / 24.0) and u[5] == 1.0 / 24.0


def test_exchange_halo_periodic():
    rng = np.random.default_rng(0)
    a = rng.random((8, 7, 6, 3))
    phys = a[2:-2, 2:-2, 2:-2].copy()
    I.exchange_halo_periodic(a, 3, 2)
    idx = [np.arange(-2, n + 2) % n for n in phys.shape[:3]]
    assert np.array_equal(a, phys[np.ix_(*idx)])
