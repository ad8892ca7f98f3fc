"""Parity of the CUDA path (through the C ABI of libdlbc_cuda.so) with the CPU oracle and with the
reference's own golden HDF5 outputs.  Tolerance everywhere (north-star):
|a-b| <= 1e-12 * max(|b|, max|b|) for populations/densities/velocities/forces; masks and the
index mapping bit-exact; total mass conserved to 1e-14 relative per step.  In EXACT arithmetic
mode with no exp() on the path the populations must be bit-identical to the oracle."""
import numpy as np
import pytest

from helpers import SC, assert_bitexact, assert_close, random_case

pytestmark = pytest.mark.gpu

cases = pytest.importorskip("dlbc_b200.cases")
from dlbc_b200 import init as I  # noqa: E402
from dlbc_b200 import lib as dlib  # noqa: E402
from dlbc_b200.connectivity import d3q19  # noqa: E402
from dlbc_b200.lattice import Lattice  # noqa: E402
from oracle import oracle as O  # noqa: E402

KERNELS = ["two_pass", "wave", "tma"]


def tma_scope(kernel, gn, mask=None, nc=2):
    """The TMA variant covers lattices without solid sites, an even nx and at most two components
    (step_tma.cuh); everything else is the register path's (AUTO picks it, an explicit request is an error)."""
    return kernel != "tma" or (gn[0] % 2 == 0 and nc <= 2 and (mask is None or not np.any(mask)))


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


# ---- the reference's in-source known-answer tests against the DEVICE site functions (lb_site.cuh) -------------
def test_site_function_kats_on_device():
    pop7 = np.zeros(19); pop7[7] = 0.1
    pop5 = np.array([0.1, 0.0, 0.1, 0.0, 0.0, 0.0, 0.0, 0.1, 0.0, 0.0, 0.1, 0.0, 0.0, 0.0, 0.1, 0.0, 0.0, 0.0, 0.0])
    # lb/density.d:50-61
    assert list(dlib.site_function("density", np.stack([pop7, pop5]))) == [0.1, 0.5]
    # lb/velocity.d:59-74 (the second with the pre-computed density 0.5), :196-217 through rawVelocity
    assert list(dlib.site_function("velocity", pop7)[0]) == [1, 1, 0]
    assert list(dlib.site_function("velocity", pop5, extra=[0.5])[0]) == [-0.2, -0.2, 0.2]
    # lb/force.d:370-384
    rho = np.array([0.0, 0.7, 1.0, 2.0])
    assert list(dlib.site_function("psi", rho, "Linear")) == [0.0, 0.7, 1.0, 2.0]
    e = dlib.site_function("psi", rho, "Exponential")
    assert e[0] == 0.0 and np.allclose(e[1:], [0.5034146962, 0.63212055882, 0.86466471676], atol=1e-9, rtol=0)
    # lb/eqdist.d:317-360: mass conserved; momentum conserved up to the shift dv (not BDist2); eqDistUnity
    rng = np.random.default_rng(7)
    pops = rng.random((20, 19))
    dv = np.full((20, 3), 0.001)
    for form in ("SecondOrder", "ThirdOrder", "BDist2"):
        eq = dlib.site_function("eqdist", pops, form, extra=dv)
        assert np.max(np.abs(dlib.site_function("density", eq) - dlib.site_function("density", pops))) < 1e-12
        if form != "BDist2":
            assert np.allclose(dlib.site_function("velocity", eq) - dv, dlib.site_function("velocity", pops), atol=1e-12)
        for i in range(20):  # the device functions against the oracle's, site by site
            assert_close(eq[i], O.site_eqdist(form, pops[i], dv[i]), what=f"eqdist {form}")
        unity = dlib.site_function("eqdist", np.eye(19)[0], form, extra=np.zeros((1, 3)))[0]
        assert abs(unity.sum() - 1.0) < 1e-15
        assert_bitexact(unity, I.eq_dist_unity(d3q19, dlib.EQ[form]), "eqDistUnity host vs device")
    for p in pops:
        assert dlib.site_function("density", p)[0] == O.site_density(p)


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
    if not tma_scope(kernel, (24, 20, 28), mask):
        pytest.skip("the TMA variant has no mask path")
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
        if not tma_scope(kernel, gn, mask):
            continue
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


# ---- the single-launch schedules in every regime the 512^3 bench runs them in -----------------------
# The wave / TMA kernels cut the lattice into y-bands of `wave_tyb` rows and x-segments of 128 / 64 sites and
# let the fused items lag `wave_la` band-planes behind the psi items; band halo rows, the ragged last band,
# the per-(band, plane, row) completion counters and the cross-band psi rewrites only exist when there is
# more than one band, and the waits only bite when more items exist than the GPU holds at once.
def _run_gpu(params, pops, mask, steps, options=None, **kw):
    p = dict(params)
    gn, nc = p.pop("gn"), p.pop("components")
    with Lattice(gn, nc, **p, **kw) as G:
        for k, v in (options or {}).items():
            G.set_option(k, v)
        if mask is not None:
            G.set_mask(mask)
        for c, a in enumerate(pops):
            G.set_populations(c, a)
        G.step(steps)
        return [G.get_populations(c) for c in range(nc)], [G.density(c) for c in range(nc)]


@pytest.mark.parametrize("solid", [0.0, 0.15])
@pytest.mark.parametrize("arith", ["exact", "fast"])
@pytest.mark.parametrize("gn", [(256, 192, 24), (200, 136, 20)])
def test_single_launch_multi_band_default_geometry(gn, arith, solid):
    """nx * ny > 16384 => several bands (256x192: 3 bands of 64 rows; 200x136: 81 + a ragged 55), 2-4
    x-segments per row, lag 4 -- the geometry of the 512^3 bench.  wave and tma must equal two_pass BIT FOR
    BIT (same site functions) and the oracle within tolerance."""
    params, pops, mask = random_case(gn, solid)
    steps = 8
    ref_f, ref_d = _run_gpu(params, pops, mask, steps, arithmetic=arith, kernel="two_pass")
    for kernel in ("wave", "tma"):
        if not tma_scope(kernel, gn, mask):
            continue
        f, d = _run_gpu(params, pops, mask, steps, arithmetic=arith, kernel=kernel)
        for c in range(2):
            assert_bitexact(f[c], ref_f[c], f"{kernel} vs two_pass {gn} {arith} solid={solid}: populations {c}")
            assert_bitexact(d[c], ref_d[c], f"{kernel} vs two_pass {gn}: density {c}")
    p = dict(params); g = p.pop("gn"); nc = p.pop("components")
    R = O.OracleLattice(g, nc, **p)
    if mask is not None:
        R.set_mask(mask)
    for c, a in enumerate(pops):
        R.set_populations(c, a)
    R.exchange_halo()
    R.step(steps)
    for c in range(2):
        assert_close(ref_f[c], R.get_populations(c), what=f"{gn} {arith} solid={solid} vs oracle: populations {c}")


@pytest.mark.parametrize("kernel", ["wave", "tma"])
@pytest.mark.parametrize("tyb,la", [(8, 2), (8, 4), (5, 3), (1, 2)])
@pytest.mark.parametrize("gn,solid", [((132, 20, 12), 0.0), ((64, 44, 10), 0.2), ((260, 19, 6), 0.0), ((72, 9, 3), 0.1),
                                      ((64, 44, 10), 0.0), ((70, 9, 3), 0.0)])
def test_single_launch_forced_bands(gn, solid, tyb, la, kernel):
    """Small lattices forced into >= 3 bands with the minimum lag (dlbc_cuda_set_option): ragged last band,
    rows that wrap around y inside a band halo, 1-3 x-segments, nz as small as 3; EXACT arithmetic,
    bit-identical to two_pass and within tolerance of the oracle."""
    params, pops, mask = random_case(gn, solid, seed=5, gwc=[-0.03, 0.04] if solid else None)
    if not tma_scope(kernel, gn, mask):
        pytest.skip("the TMA variant has no mask path")
    steps = 6
    ref_f, _ = _run_gpu(params, pops, mask, steps, arithmetic="exact", kernel="two_pass")
    f, _ = _run_gpu(params, pops, mask, steps, options=dict(wave_tyb=tyb, wave_la=la), arithmetic="exact", kernel=kernel)
    for c in range(2):
        assert_bitexact(f[c], ref_f[c], f"{kernel} tyb={tyb} la={la} {gn}: populations {c}")
    G, R = make_pair(params, pops, mask, arithmetic="exact", kernel=kernel)
    G.set_option("wave_tyb", tyb)
    G.set_option("wave_la", la)
    G.step(steps)
    R.step(steps)
    compare_state(G, R, 2, f"{kernel} tyb={tyb} la={la} {gn} vs oracle", steps=steps)
    G.close()


@pytest.mark.parametrize("kernel", ["auto", "wave", "tma"])
def test_spinodal_twin_128(kernel):
    """config 3 twin at 128^3, 10 steps against the oracle with FAST arithmetic -- what the bench
    runs -- in 8 bands of 16 rows (the default band at nx = 128 is 128 rows)."""
    params, pops, _ = cases.spinodal(128)
    G, R = make_pair(params, pops, None, kernel=kernel)
    assert G.kernel_variant in ("wave", "tma")
    G.set_option("wave_tyb", 16)
    G.step(10)
    R.step(10)
    compare_state(G, R, 2, "spinodal 128^3 t10", steps=10)
    G.close()


def test_porous_twin_64():
    """config 4 twin at 64^3 (30 % solid from Mt19937(1234)), default kernel, 4 bands, 10 steps."""
    params, pops, mask = cases.porous(64)
    G, R = make_pair(params, pops, mask)
    G.set_option("wave_tyb", 16)
    m0 = [G.total_mass(c) for c in range(2)]
    G.step(10)
    R.step(10)
    compare_state(G, R, 2, "porous 64^3 t10", steps=10)
    for c in range(2):
        assert abs(G.total_mass(c) - m0[c]) / m0[c] < 10 * 1e-14
    G.close()


def test_auto_falls_back_on_odd_nx_and_explicit_tma_fails():
    params, pops, mask = random_case((33, 6, 4))
    p = dict(params); gn = p.pop("gn"); nc = p.pop("components")
    with Lattice(gn, nc, **p) as G:
        assert G.kernel_variant == "wave"   # 16-byte TMA box starts / strides need an even nx
    with Lattice(gn, nc, **p, kernel="tma") as G:
        for c, a in enumerate(pops):
            G.set_populations(c, a)
        with pytest.raises(dlib.DlbcCudaError, match="even nx"):
            G.step(1)
    with Lattice((34, 6, 4), nc, **p, kernel="tma") as G:
        assert G.kernel_variant == "tma"
        G.set_mask((np.arange(34 * 6 * 4).reshape(34, 6, 4) % 7 == 0).astype(np.int32))
        with pytest.raises(dlib.DlbcCudaError, match="no solid sites"):
            G.step(1)                       # no mask path in the TMA variant
    with Lattice((34, 6, 4), nc, **p) as G:
        G.set_mask((np.arange(34 * 6 * 4).reshape(34, 6, 4) % 7 == 0).astype(np.int32))
        assert G.kernel_variant == "wave"   # AUTO: solid sites => register path
    with Lattice(gn, nc, **p) as G:
        with pytest.raises(dlib.DlbcCudaError):
            G.set_option("no_such_option", 1)


@pytest.mark.parametrize("gn", [(6, 5, 3), (34, 2, 2), (130, 3, 2), (2, 2, 2), (16, 16, 40), (64, 3, 3), (66, 4, 5), (126, 7, 4)])
@pytest.mark.parametrize("arith", ["exact", "fast"])
def test_tma_ragged_sizes(gn, arith):
    """the TMA variant on extents that are not multiples of its 64-site items, down to 2 (every neighbour wraps:
    both edge patches in one item), without a mask; bit-identical to two_pass, within tolerance of the oracle."""
    params, pops, _ = random_case(gn, 0.0, seed=13)
    ref_f, ref_d = _run_gpu(params, pops, None, 6, arithmetic=arith, kernel="two_pass")
    f, d = _run_gpu(params, pops, None, 6, arithmetic=arith, kernel="tma")
    for c in range(2):
        assert_bitexact(f[c], ref_f[c], f"tma vs two_pass {gn} {arith}: populations {c}")
        assert_bitexact(d[c], ref_d[c], f"tma vs two_pass {gn} {arith}: density {c}")
    G, R = make_pair(params, pops, None, arithmetic=arith, kernel="tma")
    G.step(6)
    R.step(6)
    compare_state(G, R, 2, f"tma ragged {gn} {arith} vs oracle", steps=6)
    G.close()


@pytest.mark.parametrize("kernel", KERNELS)
def test_mask_uploaded_after_populations_and_steps(kernel):
    """upload order must not matter: solid sites hold their values in BOTH population buffers
    (advection.d:83-85 copies them every step), also when the mask arrives after some steps."""
    if kernel == "tma":
        pytest.skip("the TMA variant has no mask path")
    params, pops, mask = random_case((24, 12, 10), solid=0.2, seed=9)
    p = dict(params); gn = p.pop("gn"); nc = p.pop("components")
    with Lattice(gn, nc, **p, arithmetic="exact", kernel=kernel) as G:
        for c, a in enumerate(pops):
            G.set_populations(c, a)
        G.step(3)                      # no mask yet: all sites fluid
        mid = [G.get_populations(c) for c in range(nc)]
        G.set_mask(mask)
        G.step(3)                      # an odd number: the "other" buffer is the current one now
        got = [G.get_populations(c) for c in range(nc)]
    R = O.OracleLattice(gn, nc, **p)
    for c in range(nc):
        R.set_populations(c, mid[c])
    R.set_mask(mask)
    R.exchange_halo()
    R.step(3)
    solid = mask == 1
    for c in range(nc):
        assert_bitexact(got[c][solid], mid[c][solid], f"{kernel}: solid sites keep their populations")
        assert_close(got[c], R.get_populations(c), what=f"{kernel}: mask after populations, populations {c}")
