"""The BASELINE.json workloads as (parameters, initial arrays) -- SURVEY.md section 8(d).

Each case is expressible as a reference ``.in`` file; `params` uses the reference's parameter
names (lb.tau, lb.force.gcc, ...) mapped onto the keyword arguments that both
``dlbc_b200.lattice.Lattice`` and the test oracle accept.  Random initial conditions follow the
draw order of a 1-rank reference run: one ``uniform(-p, p)`` per site of the FULL array incl.
halo in memory order (fields/init.d:68-74), a fresh Mt19937(seed_f) per fluid
(lb/init.d:110-118); the halo values are then overwritten by the first exchange (main.d:142).
"""
from __future__ import annotations

import numpy as np

from . import init as I
from .connectivity import d3q19
from .rng import Mt19937

SC2 = dict(tau=[1.0, 1.0], gcc=[[0.0, 3.6], [3.6, 0.0]], psi_form="Exponential", eqdist_form="BDist2",
           enable_shan_chen=True)


def _perturbed(gn, form, densities, perturbs, seeds):
    blk = I.Block(gn)
    pops = []
    for rho, p, seed in zip(densities, perturbs, seeds):
        a = I.init_eq_dist_perturb(blk, d3q19, form, rho, p, Mt19937(seed))
        pops.append(np.ascontiguousarray(a[blk.interior()]))
    return pops


def merging_droplets(n=64):
    """config 1: tests/runnable/tests/lb/merging-droplets-3d/merging-droplets-3d.in"""
    gn = (n, n, n)
    blk = I.Block(gn)
    pops = []
    for a, b in ((0.7, 0.04), (0.04, 0.7)):
        p = I.init_eq_dist_two_spheres_frac(blk, d3q19, I.BDIST2, a, b, 0.2, [0, 0, 0], 5.0, [0.4, 0, 0])
        pops.append(np.ascontiguousarray(p[blk.interior()]))
    return dict(gn=gn, components=2, **SC2), pops, None


def single_fluid_bgk(n=256, eqdist_form="SecondOrder"):
    """config 2: one fluid, BGK, periodic, EqDistPerturb rho = 1 +- 0.01, seed 42."""
    gn = (n, n, n) if np.isscalar(n) else tuple(n)
    params = dict(gn=gn, components=1, tau=[1.0], eqdist_form=eqdist_form, enable_shan_chen=False)
    return params, _perturbed(gn, I.EQDIST_FORMS[eqdist_form], [1.0], [0.01], [42]), None


def spinodal(n=512):
    """config 3: two-component Shan-Chen spinodal decomposition, rho = 0.37 +- 0.01 each, seeds 42, 43."""
    gn = (n, n, n) if np.isscalar(n) else tuple(n)
    return dict(gn=gn, components=2, **SC2), _perturbed(gn, I.BDIST2, [0.37, 0.37], [0.01, 0.01], [42, 43]), None


def porous_mask(gn, solid_fraction=0.3, seed=1234):
    """Solid where u < solid_fraction, u = uniform(0,1) from Mt19937(seed) over the PHYSICAL lattice in
    [x][y][z] order (handed to the reference through maskInit = File)."""
    u = Mt19937(seed).uniform(0.0, 1.0, int(np.prod(gn))).reshape(gn)
    return (u < solid_fraction).astype(np.int32)


def porous(n=512, gwc=None):
    """config 4: two fluids through a random bounce-back geometry, body force along z."""
    gn = (n, n, n) if np.isscalar(n) else tuple(n)
    mask = porous_mask(gn)
    blk = I.Block(gn, halo=0)
    pops = []
    for rho in (0.7, 0.04):
        p = I.init_eq_dist(blk, d3q19, I.BDIST2, rho)
        pops.append(I.init_eq_dist_wall(p, mask))  # fields/init.d:469-477
    params = dict(gn=gn, components=2, global_acc=[0.0, 0.0, 1e-5], **SC2)
    if gwc is not None:
        params["gwc"] = list(gwc)
    return params, pops, mask


def synthetic_layers(gn, x_count, density, perturb, form, seed):
    """x_count x-layers of a perturbed-equilibrium state drawn with numpy's PCG64 (fast): the bench's
    512^3 and larger lattices are filled layer block by layer block from this (data = synthetic)."""
    rng = np.random.default_rng(seed)
    rho = density + rng.uniform(-perturb, perturb, (x_count, gn[1], gn[2]))
    return rho[..., None] * I.eq_dist_unity(d3q19, form)
