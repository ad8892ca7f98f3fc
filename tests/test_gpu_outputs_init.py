"""SURVEY.md 8(f) rows 3 and 4 on the CUDA path: derived outputs computed on the device (pressure, corrected
velocity, the volume averages behind Laplace's law) and geometry / initial conditions generated on the device
(initMask, initEqDist*, initEqDistWall with the reference's Mt19937 draw order), each against the host-side
restatement of the reference (dlbc_b200/init.py, numpy) and the CPU oracle."""
import os

import numpy as np
import pytest

from helpers import SC, assert_bitexact, assert_close, random_case

pytestmark = pytest.mark.gpu

cases = pytest.importorskip("dlbc_b200.cases")
from dlbc_b200 import init as I  # noqa: E402
from dlbc_b200 import laplace as LP  # noqa: E402
from dlbc_b200.connectivity import d3q19  # noqa: E402
from dlbc_b200.lattice import Lattice  # noqa: E402
from dlbc_b200.rng import Mt19937  # noqa: E402
from oracle import oracle as O  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle(params, pops, mask):
    p = dict(params); gn = p.pop("gn"); nc = p.pop("components")
    R = O.OracleLattice(gn, nc, **p)
    if mask is not None:
        R.set_mask(mask)
    for c, a in enumerate(pops):
        R.set_populations(c, a)
    R.exchange_halo()
    return R


@pytest.mark.parametrize("solid", [0.0, 0.2])
def test_pressure_and_corrected_velocity(solid):
    """density.d:306-326 and velocity.d:177-185 from the oracle's densities / velocities / forces."""
    gn = (20, 12, 10)
    params, pops, mask = random_case(gn, solid, seed=3)
    p = dict(params); p.pop("gn"); nc = p.pop("components")
    R = _oracle(params, pops, mask)
    with Lattice(gn, nc, **p, arithmetic="exact") as G:
        if mask is not None:
            G.set_mask(mask)
        for c, a in enumerate(pops):
            G.set_populations(c, a)
        for steps in (0, 5):
            G.step(steps)
            R.step(steps)
            rho = [R.density(c) for c in range(nc)]
            psi = [1.0 - np.exp(-r) for r in rho]
            pr = np.zeros(gn)
            for i in range(nc):
                pr = pr + rho[i]
                for j in range(nc):
                    pr = pr + 0.5 * SC["gcc"][i][j] * psi[i] * psi[j]
            assert_close(G.pressure(), pr * (1.0 / 3.0), what=f"pressure after {steps} steps")
            for c in range(nc):
                want = R.velocity(c) + SC["tau"][c] * R.force(c) / rho[c][..., None]
                if mask is not None:
                    want[mask == 1] = 0.0
                assert_close(G.corrected_velocity(c), want, what=f"corrected velocity {c} after {steps} steps",
                             scale=1e-3, rtol=1e-12 * max(steps, 1))


def test_volume_averaged_density_and_laplace_law():
    """laplace.d:236-272 / :55-140 against numpy on the oracle's densities, t = 0 and t = 20 of a droplet."""
    gn = (24, 24, 24)
    blk = I.Block(gn)
    pops = [np.ascontiguousarray(I.init_eq_dist_sphere_frac(blk, d3q19, I.BDIST2, a, b, 0.25, [0, 0, 0], 5.0)[blk.interior()])
            for a, b in ((0.7, 0.04), (0.04, 0.7))]
    params = dict(gn=gn, components=2, **SC)
    R = _oracle(params, pops, None)
    with Lattice(gn, 2, **SC) as G:
        for c, a in enumerate(pops):
            G.set_populations(c, a)
        for steps in (0, 20):
            G.step(steps)
            R.step(steps)
            rho = [R.density(c) for c in range(2)]
            for off in ([12, 12, 12], [1, 1, 1], [0, 5, 23]):  # the last one hangs over the edge: not wrapped
                sl = tuple(slice(max(o - 1, 0), min(o + 1, n - 1) + 1) for o, n in zip(off, gn))
                want = [float(np.mean(r[sl])) for r in rho]
                got = G.volume_averaged_density(off, 1)
                assert_close(got, want, what=f"volume average at {off}")
            m = LP.measure(G, gn, SC["gcc"])
            in_den = [float(np.mean(r[11:14, 11:14, 11:14])) for r in rho]
            out_den = [float(np.mean(r[0:3, 0:3, 0:3])) for r in rho]
            assert_close(m["inPres"], LP.pressure(in_den, SC["gcc"]), what="inPres")
            assert_close(m["outPres"], LP.pressure(out_den, SC["gcc"]), what="outPres")
        assert np.isfinite(m["sigma"]) and 3.0 < m["measuredR"] < 9.0  # (the transient sigma of a 24^3 droplet can have either sign)


def test_laplace_droplet_3d_asc_golden():
    """The reference's long-run physics KAT: tests/runnable/tests/lb/laplace-droplet-3d, laplace-*.asc rows at
    t = 0, 100, ..., 500 (printed with 7 significant digits), replayed through the CUDA path."""
    from dlbc_b200 import params as P
    from dlbc_b200 import replay as RP
    golden = np.loadtxt(os.path.join(ROOT, "tests", "golden", "laplace-droplet-3d.asc"), comments="#")
    p = P.parse_file(os.path.join(ROOT, "tests", "golden", "in", "laplace-droplet-3d.in"), [])
    gn, C, kw = RP.lattice_kwargs(p)
    blk = I.Block(gn)
    pops = RP.initial_populations(p, blk, RP.initial_mask(p, blk))
    with Lattice(gn, C, **kw) as G:
        for c, a in enumerate(pops):
            G.set_populations(c, a, halo=blk.halo)
        prev = float("nan")
        for row in golden[:6]:
            t = int(row[2])
            G.step(t - G.timestep)
            m = LP.measure(G, gn, kw["gcc"], kw["psi_form"], previous_sigma=prev)
            prev = m["sigma"]
            for name, col in (("sigma", 3), ("measuredR", 5), ("inPres", 6), ("outPres", 7)):
                assert abs(m[name] - row[col]) <= 2e-6 * max(abs(row[col]), 1e-2), (t, name, m[name], row[col])
            if t >= 200:
                assert abs(m["rel"] - row[4]) <= 1e-4, (t, m["rel"], row[4])


# ---- device-side geometry and initial conditions -------------------------------------------------------------
@pytest.mark.parametrize("kind,kw", [("None", {}), ("Box", {}), ("Tube", dict(axis=2)), ("Tube", dict(axis=0)),
                                     ("Plates", dict(axis=1, wall_offset=2)),
                                     ("Walls", dict(thickness=2, dimension_has_wall=[[True, False], [False, False], [True, True]]))])
def test_init_mask_on_device(kind, kw):
    gn = (14, 10, 12)
    blk = I.Block(gn, halo=0)
    want = {"None": lambda: I.mask_none(blk), "Box": lambda: I.mask_box(blk), "Tube": lambda: I.mask_tube(blk, kw.get("axis", 0)),
            "Plates": lambda: I.mask_plates(blk, kw.get("axis", 0), kw.get("wall_offset", 0)),
            "Walls": lambda: I.mask_walls(blk, kw.get("dimension_has_wall"), kw.get("thickness", 1))}[kind]()
    with Lattice(gn, 1, tau=[1.0]) as G:
        G.init_mask(kind, **kw)
        assert_bitexact(G.get_mask(), want.astype(np.int32), f"initMask {kind}")


@pytest.mark.parametrize("gn", [(12, 10, 8), (64, 64, 64), (30, 7, 50)])
def test_init_populations_on_device_bitexact(gn):
    """EqDist / EqDistRandom / EqDistPerturb / EqDistPerturbFrac + initEqDistWall generated on the device equal the
    host restatement bit for bit (Mt19937 stream, draw order over the full array incl. halo, Phobos uniform)."""
    blk = I.Block(gn)
    mask = I.mask_plates(I.Block(gn, halo=0), 2, 1)
    with Lattice(gn, 3, tau=[1.0] * 3, eqdist_form="BDist2") as G:
        G.init_mask("Plates", axis=2, wall_offset=1)
        # three fluids: a reseeding perturbed one, then two drawing from the shared global generator one after the other
        G.init_populations(0, "EqDistPerturb", 0.37, perturb=0.01, seed=42)
        nfull = int(np.prod(blk.shape_h))
        G.init_populations(1, "EqDistRandom", 0.5, seed=7)
        G.init_populations(2, "EqDistPerturbFrac", 0.8, perturb=0.05, seed=7, skip=nfull)
        G.init_wall()
        shared = Mt19937(7)
        want = [I.init_eq_dist_perturb(blk, d3q19, I.BDIST2, 0.37, 0.01, Mt19937(42)),
                I.init_eq_dist_random(blk, d3q19, I.BDIST2, 0.5, shared),
                I.init_eq_dist_perturb_frac(blk, d3q19, I.BDIST2, 0.8, 0.05, shared)]
        for c in range(3):
            w = I.init_eq_dist_wall(np.ascontiguousarray(want[c][blk.interior()]), mask)
            assert_bitexact(G.get_populations(c), w, f"device init of fluid {c} on {gn}")
    with Lattice(gn, 1, tau=[1.0], eqdist_form="ThirdOrder") as G:
        G.init_populations(0, "EqDist", 0.9)
        w = I.init_eq_dist(I.Block(gn, halo=0), d3q19, I.THIRD_ORDER, 0.9)
        assert_bitexact(G.get_populations(0), w, "device EqDist")


def test_init_profiles_on_device_bitexact():
    """spheres, two spheres, cylinder and lamellae (fields/init.d:84-313) generated on the device == host restatement."""
    gn = (20, 16, 24)
    blk = I.Block(gn)
    it = blk.interior()
    with Lattice(gn, 4, tau=[1.0] * 4, eqdist_form="BDist2") as G:
        G.init_spheres(0, 0.7, 0.04, 0.25, 5.0, frac=True)                                               # laplace-droplet-3d
        G.init_spheres(1, 0.04, 0.7, 0.2, 5.0, separation=[0.4, 0, 0], frac=True)                       # merging-droplets-3d
        G.init_spheres(2, 0.5, 0.1, 4.5, 0.0, offset=[1.0, -2.0, 0.5], cylinder_axis=1)                 # sharp cylinder
        G.init_lamellae(3, [0.7, 0.04, 0.3], [0.25, 0.5, 0.25], 2, 3.0, frac=True)                      # bdist2-sc-3d style
        want = [I.init_eq_dist_sphere_frac(blk, d3q19, I.BDIST2, 0.7, 0.04, 0.25, [0, 0, 0], 5.0),
                I.init_eq_dist_two_spheres_frac(blk, d3q19, I.BDIST2, 0.04, 0.7, 0.2, [0, 0, 0], 5.0, [0.4, 0, 0]),
                I.init_eq_dist_cylinder(blk, d3q19, I.BDIST2, 0.5, 0.1, 1, 4.5, [1.0, -2.0, 0.5], 0.0),
                I.init_eq_dist_lamellae_frac(blk, d3q19, I.BDIST2, [0.7, 0.04, 0.3], [0.25, 0.5, 0.25], 2, 3.0)]
        for c in range(4):
            assert_bitexact(G.get_populations(c), np.ascontiguousarray(want[c][it]), f"device profile init {c}")


def test_spinodal_from_device_init_matches_host_init():
    """BASELINE configs[2] twin started from device-generated populations steps exactly like the host-initialised one."""
    params, pops, _ = cases.spinodal(32)
    p = dict(params); gn = p.pop("gn"); nc = p.pop("components")
    with Lattice(gn, nc, **p, arithmetic="exact") as A, Lattice(gn, nc, **p, arithmetic="exact") as B:
        for c, a in enumerate(pops):
            A.set_populations(c, a)
        B.init_mask("None")
        for c, seed in enumerate((42, 43)):
            B.init_populations(c, "EqDistPerturb", 0.37, perturb=0.01, seed=seed)
        B.init_wall()
        A.step(10)
        B.step(10)
        for c in range(nc):
            assert_bitexact(A.get_populations(c), B.get_populations(c), f"populations {c} after 10 steps")
