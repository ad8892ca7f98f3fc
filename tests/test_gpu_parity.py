"""Parity of the CUDA path (through the C ABI of libdlbc_cuda.so) with the CPU oracle and with the
reference's own golden HDF5 outputs.  Tolerance everywhere (north-star):
|a-b| <= 1e-12 * max(|b|, max|b|) for populations/densities/velocities/forces; masks and the
index mapping bit-exact; total mass conserved to 1e-14 relative per step.  In EXACT arithmetic
mode with no exp() on the path the populations must be bit-identical to the oracle."""
import numpy as np
import pytest

from helpers import SC, assert_bitexact, assert_close

pytestmark = pytest.mark.gpu

cases = pytest.importorskip("dlbc_b200.cases")
from dlbc_b200 import init as I  # noqa: E402
from dlbc_b200 import lib as dlib  # noqa: E402
from dlbc_b200.connectivity import d3q19  # noqa: E402
from dlbc_b200.lattice import Lattice  # noqa: E402
from oracle import oracle as O  # noqa: E402

KERNELS = ["two_pass", "zmarch", "zpipe", "wave"]


def make_pair(params, pops, mask, **gpu_kw):
    """The same case on the GPU and in the oracle."""
    p = dict(params)
    gn = p.pop("gn")
    nc = p.pop("components")
    G = Lattice(gn, nc, **p, **gpu_kw)
    R = O.OracleLattice(gn, nc, **p)
    if mask is not None:
        G.set_mask(mask)
        R.set_mask(mask)
    for c, a in enumerate(pops):
        G.set_populations(c, a)
        R.set_populations(c, a)
    R.exchange_halo()  # main.d:142
    return G, R


def compare_force(G, R, nc, gcc, what):
    """The Shan-Chen force is a sum of 18 * C terms w_q * psi1 * psi2 * g of alternating sign; the 1-ulp
    freedom of exp() in psi moves each term by ~1e-16 of ITS size, so the tolerance is relative to
    that term scale (a near-uniform state has |F| ~ 1e-7 from terms ~ 1e-2)."""
    psimax = max(float(np.max(np.abs(R.psi(c)))) for c in range(nc))
    term = max(abs(x) for row in gcc for x in row) * psimax * psimax
    for c in range(nc):
        assert_close(G.force(c), R.force(c), what=f"{what}: force {c}", scale=term)


def compare_state(G, R, nc, what, bitexact=False, steps=1):
    """Populations and densities: 1e-12 relative to the field scale, whatever the step count.
    Velocities are differences of populations (|u| ~ 1e-5 from terms ~ 1e-1), so their round-off
    is measured against a much smaller scale: the north-star's tolerance is 1e-12 relative PER
    STEP, i.e. 1e-12 * steps after `steps` steps."""
    for c in range(nc):
        a, b = G.get_populations(c), R.get_populations(c)
        if bitexact:
            assert_bitexact(a, b, f"{what}: populations {c}")
        else:
            assert_close(a, b, what=f"{what}: populations {c}")
        assert_close(G.density(c), R.density(c), what=f"{what}: density {c}")
        if bitexact:
            assert_bitexact(G.velocity(c), R.velocity(c), f"{what}: velocity {c}")
        else:
            assert_close(G.velocity(c), R.velocity(c), rtol=1e-12 * steps, what=f"{what}: velocity {c}")


def test_tables_bitexact():
    v, b, w, css = dlib.tables()
    rv, rb, rw, rcss = O.tables(3, 19)
    assert (v == rv).all() and (b == rb).all()
    assert_bitexact(w, rw, "weights")
    assert css == rcss
    assert (v == d3q19.velocities).all()


def test_upload_download_roundtrip_bitexact():
    rng = np.random.default_rng(1)
    gn = (20, 12, 9)
    with Lattice(gn, 2) as G:
        a = rng.random((24, 16, 13, 19))            # with the reference's halo of 2
        b = rng.random(gn + (19,))
        G.set_populations(0, a, halo=2)
        G.set_populations(1, b)
        assert_bitexact(G.get_populations(0), a[2:-2, 2:-2, 2:-2], "halo-2 upload")
        out = G.get_populations(1, halo=2)
        assert_bitexact(out[2:-2, 2:-2, 2:-2], b, "halo-2 download")
        assert (out[:2] == 0).all() and (out[:, :2] == 0).all() and (out[..., :2, :] == 0).all()
        # x-range streaming
        G.set_populations_xrange(1, a[5:12], 3, halo=2)
        got = G.get_populations_xrange(1, 3, 7)
        assert_bitexact(got, a[5:12, 2:-2, 2:-2], "xrange")
        assert_bitexact(G.get_populations(1)[:3], b[:3], "xrange left the other layers alone")
        m = (rng.random((24, 16, 13)) < 0.3).astype(np.int32)
        G.set_mask(m, halo=2)
        assert_bitexact(G.get_mask(), m[2:-2, 2:-2, 2:-2], "mask")


def test_argument_errors():
    with pytest.raises(dlib.DlbcCudaError):
        Lattice((8, 8, 8), 1, device=99)
    with pytest.raises(ValueError):
        Lattice((8, 8, 9), 1, rank=0, nranks=2, nccl_id=bytes(128))   # lattice.d:233-241
    with Lattice((8, 8, 8), 1) as G:
        with pytest.raises(dlib.DlbcCudaError, match="0 .None. or 1"):
            G.set_mask(np.full((8, 8, 8), 2, dtype=np.int32))
        with pytest.raises(dlib.DlbcCudaError):
            G.density(3)
        with pytest.raises(ValueError):
            G.set_populations(0, np.zeros((8, 8, 8, 18)))


# ---- the reference's golden files, replayed on the GPU -----------------------------------------
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("arith", ["exact", "fast"])
def test_golden_checkpoint_3d_restart(golden, arith, kernel):
    """technical/checkpoint-3d: restore populations at t=0, 5 steps, compare populations (16^3)."""
    g = golden("checkpoint-3d")
    n = g["cp-fluids0-t0"].shape[0]
    with Lattice((n, n, n), 2, **SC, arithmetic=arith, kernel=kernel) as G:
        G.set_populations(0, g["cp-fluids0-t0"])
        G.set_populations(1, g["cp-fluids1-t0"])
        G.set_mask(g["cp-mask-t0"])
        G.step(5)
        for c, name in enumerate(("red", "blue")):
            assert_close(G.get_populations(c), g[f"population-{name}-t5"], what=f"{name} t5")
            assert_close(G.get_populations(c), g[f"cp-fluids{c}-t5"], what=f"cp {name} t5")
            assert_close(G.density(c), g[f"density-{name}-t5"], what=f"density {name} t5")
        assert_close(G.colour(0, 1), g["colour-red-blue-t5"], what="colour t5")
        assert_bitexact(G.get_mask(), g["cp-mask-t5"], "mask")


@pytest.mark.parametrize("kernel", KERNELS)
def test_golden_merging_droplets_3d(golden, kernel):
    """BASELINE config 1 (64^3, TwoSpheresFrac, SC exponential, BDist2): goldens at t=0 and t=10."""
    g = golden("merging-droplets-3d")
    params, pops, _ = cases.merging_droplets()
    G, R = make_pair(params, pops, None, kernel=kernel)
    assert_bitexact(G.density(0), g["density-red-t0"], "t0 red")
    assert_bitexact(G.density(1), g["density-blue-t0"], "t0 blue")
    m0 = [G.total_mass(c) for c in range(2)]
    G.step(10)
    R.step(10)
    assert_close(G.density(0), g["density-red-t10"], what="t10 red vs golden")
    assert_close(G.density(1), g["density-blue-t10"], what="t10 blue vs golden")
    assert_close(G.colour(0, 1), g["density-red-t10"] - g["density-blue-t10"], what="t10 colour vs golden")
    compare_state(G, R, 2, "merging droplets t10 vs oracle", steps=10)
    compare_force(G, R, 2, SC["gcc"], "merging droplets")
    for c in range(2):
        assert abs(G.total_mass(c) - m0[c]) / m0[c] < 10 * 1e-14
    G.close()


def test_golden_merging_droplets_3d_t1000(golden):
    g = golden("merging-droplets-3d")
    params, pops, _ = cases.merging_droplets()
    p = dict(params); gn = p.pop("gn"); nc = p.pop("components")
    with Lattice(gn, nc, **p) as G:
        for c, a in enumerate(pops):
            G.set_populations(c, a)
        G.step(1000)
        # the reference's own tolerance for this file is 1e-14 absolute (merging-droplets-3d.json)
        assert np.max(np.abs(G.density(0) - g["density-red-t1000"])) <= 1e-12
        assert np.max(np.abs(G.density(1) - g["density-blue-t1000"])) <= 1e-12


@pytest.mark.parametrize("kernel", KERNELS)
def test_golden_bdist2_sc_3d_populations(golden, kernel):
    """lb/bdist2-sc-3d: full 19 populations at t=10 (Lamellae init, SC)."""
    g = golden("bdist2-sc-3d")
    n = 64
    blk = I.Block((n, n, n))
    with Lattice((n, n, n), 2, **SC, kernel=kernel) as G:
        for c, v in enumerate(((0.7, 0.04), (0.04, 0.7))):
            p = I.init_eq_dist_lamellae(blk, d3q19, I.BDIST2, v, [32, 32], 2, 0.0)
            G.set_populations(c, p, halo=2)
        for t, steps in ((10, 10), (100, 90)):
            G.step(steps)
            for c, name in enumerate(("red", "blue")):
                pop = G.get_populations(c)
                # the golden file holds the line along z (the run is uniform in x and y)
                assert np.array_equal(pop, np.broadcast_to(pop[0:1, 0:1], pop.shape))
                assert_close(pop[0, 0], g[f"population-{name}-t{t}"], what=f"population {name} t{t}")
                assert_close(G.density(c)[0, 0], g[f"density-{name}-t{t}"], what=f"density {name} t{t}")
            assert_close(G.colour(0, 1)[0, 0], g[f"colour-red-blue-t{t}"], what=f"colour t{t}")


def test_golden_bdist2_3d(golden):
    """lb/bdist2-3d: pull + BDist2 without Shan-Chen on 4^3 (smaller than a warp, all wrap-around)."""
    g = golden("bdist2-3d")
    with Lattice((4, 4, 4), 2, tau=[1.0, 1.0], eqdist_form="BDist2", arithmetic="exact") as G:
        for c, name in enumerate(("red", "blue")):
            G.set_populations(c, g[f"population-{name}-t0"])
        G.step(10)
        for c, name in enumerate(("red", "blue")):
            # the reference compares this case bit for bit; EXACT arithmetic has no exp() here
            assert_bitexact(G.get_populations(c), g[f"population-{name}-t10"], f"population {name} t10")
            assert_bitexact(G.density(c), g[f"density-{name}-t10"], f"density {name} t10")


# ---- against the oracle on seeded synthetic inputs -----------------------------------------------
@pytest.mark.parametrize("form", ["SecondOrder", "ThirdOrder", "BDist2"])
def test_single_fluid_bgk_exact_is_bit_identical(form):
    """config 2 twin (32^3): no exp() on the path => EXACT mode reproduces the oracle bit for bit."""
    params, pops, _ = cases.single_fluid_bgk(32, form)
    params["global_acc"] = [1e-6, -2e-6, 3e-6]
    params["tau"] = [0.8]
    G, R = make_pair(params, pops, None, arithmetic="exact")
    G.step(20)
    R.step(20)
    compare_state(G, R, 1, f"bgk {form} exact", bitexact=True)
    G.close()


def test_single_fluid_bgk_fast_100_steps():
    params, pops, _ = cases.single_fluid_bgk(32)
    G, R = make_pair(params, pops, None, arithmetic="fast")
    m0 = G.total_mass(0)
    G.step(100)
    R.step(100)
    compare_state(G, R, 1, "bgk fast t100", steps=100)
    assert abs(G.total_mass(0) - m0) / m0 < 100 * 1e-14
    G.close()


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("arith", ["exact", "fast"])
def test_spinodal_twin(arith, kernel):
    """config 3 twin (32^3, 50 steps)."""
    params, pops, _ = cases.spinodal(32)
    G, R = make_pair(params, pops, None, arithmetic=arith, kernel=kernel)
    G.step(50)
    R.step(50)
    compare_state(G, R, 2, f"spinodal {arith}", steps=50)
    G.close()


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("gwc", [None, [-0.05, 0.05]])
def test_porous_twin(gwc, kernel):
    """config 4 twin (24x20x28, 30 % random solid): bounce-back + SC next to walls + wall adhesion."""
    params, pops, mask = cases.porous((24, 20, 28), gwc=gwc)
    G, R = make_pair(params, pops, mask, kernel=kernel)
    assert_bitexact(G.get_mask(), mask, "mask round trip")
    m0 = [G.total_mass(c) for c in range(2)]
    G.step(20)
    R.step(20)
    compare_state(G, R, 2, f"porous gwc={gwc}", steps=20)
    solid = mask == 1
    compare_force(G, R, 2, SC["gcc"], f"porous gwc={gwc}")
    for c in range(2):
        assert_bitexact(G.get_populations(c)[solid], pops[c][solid], "solid sites never change")
        assert (G.velocity(c)[solid] == 0).all() and (G.colour(0, 1)[solid] == 0).all()
        assert abs(G.total_mass(c) - m0[c]) / m0[c] < 20 * 1e-14
        assert abs(G.total_mass(c) - R.total_mass(c)) / m0[c] < 1e-12
    assert_bitexact(G.get_mask(), mask, "mask unchanged by stepping")
    G.close()


@pytest.mark.parametrize("gn", [(7, 5, 3), (33, 2, 2), (130, 3, 2), (2, 2, 2), (16, 16, 40)])
def test_ragged_sizes(gn):
    """extents that are not multiples of the block/warp size, down to 2 (every neighbour wraps).
    Extents of 1 are outside the reference's domain: its halo of 2 is filled from the 2 outermost
    PHYSICAL layers (fields/parallel.d:58-68), which do not exist for n = 1."""
    rng = np.random.default_rng(7)
    mask = (rng.random(gn) < 0.2).astype(np.int32)
    pops = []
    for rho in (0.6, 0.1):
        a = (rho + 0.01 * rng.random(gn))[..., None] * I.eq_dist_unity(d3q19, I.BDIST2)
        pops.append(I.init_eq_dist_wall(a, mask))
    params = dict(gn=gn, components=2, global_acc=[1e-5, 0, 0], **SC)
    for kernel in KERNELS:
        G, R = make_pair(params, pops, mask, kernel=kernel)
        G.step(6)
        R.step(6)
        compare_state(G, R, 2, f"ragged {gn} {kernel}", steps=6)
        G.close()


def test_three_components_linear_psi_third_order():
    """paths no reference golden pins: C = 3, Linear psi, ThirdOrder, asymmetric gcc, unequal tau."""
    gn = (12, 10, 14)
    rng = np.random.default_rng(3)
    pops = [(r + 0.02 * rng.random(gn))[..., None] * I.eq_dist_unity(d3q19, I.THIRD_ORDER) for r in (0.5, 0.3, 0.2)]
    params = dict(gn=gn, components=3, tau=[1.0, 0.7, 1.3], gcc=[[0.0, 0.4, 0.2], [0.4, 0.1, 0.3], [0.25, 0.3, 0.0]],
                  psi_form="Linear", eqdist_form="ThirdOrder", enable_shan_chen=True, global_acc=[0, 1e-6, 0])
    for arith in ("exact", "fast"):
        G, R = make_pair(params, pops, None, arithmetic=arith)
        G.step(10)
        R.step(10)
        compare_state(G, R, 3, f"C=3 linear {arith}", bitexact=(arith == "exact"), steps=10)
        compare_force(G, R, 3, params["gcc"], f"C=3 {arith}")
        G.close()


def test_four_components():
    gn = (10, 8, 6)
    rng = np.random.default_rng(4)
    pops = [(0.25 + 0.02 * rng.random(gn))[..., None] * I.eq_dist_unity(d3q19, I.BDIST2) for _ in range(4)]
    gcc = [[0.0 if i == j else 0.9 for j in range(4)] for i in range(4)]
    params = dict(gn=gn, components=4, tau=[1.0] * 4, gcc=gcc, psi_form="Exponential", eqdist_form="BDist2",
                  enable_shan_chen=True)
    G, R = make_pair(params, pops, None)
    G.step(8)
    R.step(8)
    compare_state(G, R, 4, "C=4", steps=8)
    G.close()


# ---- full-size properties (sizes the oracle cannot replay in seconds) ----------------------------
def test_full_size_mass_and_linearity_256():
    """config 2 at its full 256^3: mass conserved to round-off per step; momentum conserved without
    forcing; translation invariance (rolling the input by (3,5,7) rolls the output) pins the
    periodic index arithmetic at a size the oracle cannot replay quickly."""
    n = 256
    gn = (n, n, n)
    layers = cases.synthetic_layers(gn, 16, 1.0, 0.01, I.SECOND_ORDER, 5)
    with Lattice(gn, 1, tau=[1.0]) as G:
        for x0 in range(0, n, 16):
            G.set_populations_xrange(0, np.roll(layers, x0 // 16, axis=1), x0)
        m0, p0 = G.total_mass(0), G.total_momentum(0)
        G.step(10)
        m1, p1 = G.total_mass(0), G.total_momentum(0)
        assert abs(m1 - m0) / m0 < 10 * 1e-14
        assert np.all(np.abs(p1 - p0) < 1e-12 * m0)
        probe = G.get_populations_xrange(0, 100, 2)
    with Lattice(gn, 1, tau=[1.0]) as G:
        for x0 in range(0, n, 16):
            blk = np.roll(layers, x0 // 16, axis=1)
            G.set_populations_xrange(0, np.roll(blk, (5, 7), axis=(1, 2)), (x0 + 16 * 3) % n)
        G.step(10)
        shifted = G.get_populations_xrange(0, (100 + 48) % n, 2)
    assert_bitexact(shifted, np.roll(probe, (5, 7), axis=(1, 2)), "translation invariance")
