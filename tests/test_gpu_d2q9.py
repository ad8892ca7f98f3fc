"""SURVEY.md 8(f) row 4, second half: the reference's 2-D lattice (Lattice!(d2q9)) on the device (csrc/d2q9.cu), through
the same C ABI (a context created with d = 2, q = 9), against the CPU oracle's D2Q9 instantiation -- bit for bit with
linear psi (every operation in the reference's order), to the parity tolerance with exp() in psi -- and against the
reference's own 2-D golden files (lb/poiseuille-2d)."""
import numpy as np
import pytest

from helpers import assert_bitexact, assert_close

pytestmark = pytest.mark.gpu

from dlbc_b200 import init as I  # noqa: E402
from dlbc_b200.connectivity import d2q9  # noqa: E402
from dlbc_b200.lattice import Lattice  # noqa: E402
from dlbc_b200.lib import DlbcCudaError  # noqa: E402
from oracle import oracle as O  # noqa: E402


def _case(gn, ncomp, solid, seed):
    """Random positive populations (no particular equilibrium: every term of every formula is exercised) + mask."""
    rng = np.random.default_rng(seed)
    mask = (rng.random(gn) < solid).astype(np.int32) if solid > 0 else None
    pops = []
    for c in range(ncomp):
        rho = (0.1 + 0.6 / (c + 1)) + 0.05 * rng.random(gn)
        a = rho[..., None] * I.eq_dist_unity(d2q9, I.SECOND_ORDER) * (1.0 + 0.02 * rng.random(gn + (9,)))
        pops.append(I.init_eq_dist_wall(a, mask) if mask is not None else a)
    return pops, mask


def _pair(gn, ncomp, pops, mask, **kw):
    R = O.OracleLattice(gn, ncomp, **kw)
    G = Lattice(gn, ncomp, **kw, arithmetic="exact")
    for L in (R, G):
        if mask is not None:
            L.set_mask(mask)
        for c, a in enumerate(pops):
            L.set_populations(c, a)
    R.exchange_halo()
    return R, G


def _compare(R, G, ncomp, exact, what):
    cmp = assert_bitexact if exact else (lambda a, b, w: assert_close(a, b, what=w, scale=1e-3))
    for c in range(ncomp):
        cmp(G.get_populations(c), R.get_populations(c), f"{what}: populations {c}")
        cmp(G.density(c), R.density(c), f"{what}: density {c}")
        cmp(G.velocity(c), R.velocity(c), f"{what}: velocity {c}")
        cmp(G.force(c), R.force(c), f"{what}: force {c}")
    if ncomp > 1:
        cmp(G.colour(0, 1), R.colour(0, 1), f"{what}: colour")


GCC2 = [[0.0, 3.6], [3.6, 0.0]]


@pytest.mark.parametrize("eq", ["SecondOrder", "ThirdOrder", "BDist2"])
@pytest.mark.parametrize("solid", [0.0, 0.2])
def test_d2q9_two_components_linear_psi_bit_exact(eq, solid):
    gn = (37, 22)
    pops, mask = _case(gn, 2, solid, seed=3)
    kw = dict(tau=[1.0, 0.8], gcc=GCC2, gwc=[0.05, -0.03] if solid else None, global_acc=[1e-5, -2e-5], psi_form="Linear",
              eqdist_form=eq, enable_shan_chen=True)
    R, G = _pair(gn, 2, pops, mask, **kw)
    with G:
        _compare(R, G, 2, True, f"{eq} t0")
        for steps in (1, 7):
            R.step(steps); G.step(steps)
            _compare(R, G, 2, True, f"{eq} solid {solid} after {steps} more steps")
        if mask is not None:
            assert_bitexact(G.get_mask(), mask, "mask round trip")
            for c in range(2):  # solid sites never change (advection.d:83-85)
                assert_bitexact(G.get_populations(c)[mask == 1], pops[c][mask == 1], "solid sites")
        assert G.timestep == 8


@pytest.mark.parametrize("eq", ["SecondOrder", "BDist2"])
def test_d2q9_exponential_psi(eq):
    gn = (32, 48)
    pops, mask = _case(gn, 2, 0.1, seed=5)
    kw = dict(tau=[1.0, 1.0], gcc=GCC2, global_acc=[0.0, 1e-5], psi_form="Exponential", eqdist_form=eq, enable_shan_chen=True)
    R, G = _pair(gn, 2, pops, mask, **kw)
    with G:
        R.step(20); G.step(20)
        _compare(R, G, 2, False, f"{eq} exponential psi")
        for c in range(2):
            assert abs(G.total_mass(c) - R.total_mass(c)) <= 1e-12 * R.total_mass(c)
            assert_close(G.total_momentum(c)[:2], R.total_momentum(c)[:2], what="momentum", scale=1.0)
            assert G.total_momentum(c)[2] == 0.0


@pytest.mark.parametrize("ncomp,gn", [(1, (16, 16)), (3, (9, 14)), (4, (5, 6)), (2, (2, 8)), (2, (33, 2)), (1, (130, 3))])
def test_d2q9_component_counts_and_ragged_sizes(ncomp, gn):
    # (extents below the halo width of 2 are outside what the reference can run: its halo exchange copies 2 layers)
    pops, mask = _case(gn, ncomp, 0.15 if min(gn) > 2 else 0.0, seed=ncomp)
    gcc = [[0.0 if i == j else 1.5 + 0.1 * (i + j) for j in range(ncomp)] for i in range(ncomp)]
    kw = dict(tau=[1.0 - 0.05 * c for c in range(ncomp)], gcc=gcc, global_acc=[2e-5, 1e-5], psi_form="Linear",
              eqdist_form="ThirdOrder", enable_shan_chen=ncomp > 1)
    R, G = _pair(gn, ncomp, pops, mask, **kw)
    with G:
        R.step(5); G.step(5)
        _compare(R, G, ncomp, True, f"C = {ncomp}, {gn}")


def test_d2q9_halo_2_host_arrays():
    """Field.arr as the D host hands it over: two halo layers around the physical sites; only those are moved."""
    gn = (12, 10)
    pops, mask = _case(gn, 2, 0.2, seed=9)
    with Lattice(gn, 2, tau=[1.0, 1.0], arithmetic="exact") as G:
        big = [np.full((16, 14, 9), np.nan) for _ in pops]
        for b, a in zip(big, pops):
            b[2:-2, 2:-2] = a
        bm = np.zeros((16, 14), dtype=np.int32)
        bm[2:-2, 2:-2] = mask
        G.set_mask(bm, halo=2)
        for c, b in enumerate(big):
            G.set_populations(c, b, halo=2)
        for c in range(2):
            assert_bitexact(G.get_populations(c), pops[c], "halo-2 upload")
            back = G.get_populations(c, halo=2)
            assert_bitexact(back[2:-2, 2:-2], pops[c], "halo-2 download")
            assert (back[:2] == 0).all() and (back[:, -2:] == 0).all()
        assert_bitexact(G.get_mask(halo=2)[2:-2, 2:-2], mask, "mask")
        with pytest.raises(DlbcCudaError):
            G.set_mask(np.full(gn, 2, dtype=np.int32))


def test_d2q9_poiseuille_golden(golden):
    """lb/poiseuille-2d (reference-data): plates, bounce-back, body force, t = 0 and t = 1000."""
    g = golden("poiseuille-2d")
    blk = I.Block((32, 4))
    mask = I.mask_plates(blk, 0, 0)
    with Lattice((32, 4), 2, tau=[1.0, 1.0], global_acc=[0.0, 0.0001], eqdist_form="SecondOrder", arithmetic="exact") as L:
        L.set_mask(mask, halo=2)
        for c, rho in enumerate((0.7, 1.0)):
            p = I.init_eq_dist(blk, d2q9, I.SECOND_ORDER, rho)
            L.set_populations(c, I.init_eq_dist_wall(p, mask), halo=2)
        assert_bitexact(L.get_mask(), g["mask-t0"], "mask")
        for c, name in enumerate(("red", "blue")):
            assert_bitexact(L.density(c), g[f"density-{name}-t0"], "density t0")
            assert_bitexact(L.velocity(c), g[f"velocity-{name}-t0"], "velocity t0")
        L.step(1000)
        for c, name in enumerate(("red", "blue")):
            assert np.max(np.abs(L.density(c) - g[f"density-{name}-t1000"])) <= 5e-15
            assert np.max(np.abs(L.velocity(c) - g[f"velocity-{name}-t1000"])) <= 5e-15
            assert np.max(np.abs(L.force(c) - g[f"force-{name}-t1000"])) == 0.0
            assert (L.velocity(c)[[0, -1]] == 0).all()
            assert_bitexact(L.get_populations(c)[[0, -1]], np.broadcast_to(np.eye(9)[0], (2, 4, 9)).copy())
        assert_bitexact(L.get_mask(), g["mask-t1000"], "mask unchanged")


def test_d2q9_context_rejects_3d_only_entry_points():
    with Lattice((8, 8), 1, arithmetic="exact") as G:
        for call in (G.pressure, lambda: G.init_mask("None"), lambda: G.volume_averaged_density((4, 4, 0))):
            with pytest.raises(DlbcCudaError, match="D2Q9"):
                call()
        assert G.kernel_variant == "two_pass" and G.launch_count >= 0 and G.device_bytes > 0
    with pytest.raises(DlbcCudaError):
        Lattice((8, 8), 1, nranks=1, tau=[-1.0])
