"""Shared helpers for the parity tests."""
import numpy as np

# |a-b| <= 1e-12 * max(|b|, s), s = max-abs of the golden field
RTOL = 1e-12


def assert_close(a, b, rtol=RTOL, what="", scale=None):
    """`scale`: natural magnitude of the terms the field is summed from, for fields that are small
    only through cancellation (forces of a near-uniform state, velocities of a fluid at rest)."""
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    s = float(np.max(np.abs(b))) if b.size else 0.0
    if scale is not None:
        s = max(s, float(scale))
    bound = rtol * np.maximum(np.abs(b), s)
    err = np.abs(a - b)
    bad = err > bound
    assert not bad.any(), (
        f"{what}: {int(bad.sum())} of {bad.size} values outside {rtol:g} relative; "
        f"max abs err {float(err.max()):.3e} (field scale {s:.3e})")


def assert_bitexact(a, b, what=""):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape and a.dtype == b.dtype, (what, a.shape, b.shape, a.dtype, b.dtype)
    if a.dtype.kind == "f":
        same = a.view(np.uint64) == b.view(np.uint64)
    else:
        same = a == b
    assert same.all(), f"{what}: {int((~same).sum())} of {same.size} values differ bitwise"


SC = dict(tau=[1.0, 1.0], gcc=[[0.0, 3.6], [3.6, 0.0]], psi_form="Exponential",
          eqdist_form="BDist2", enable_shan_chen=True)


def random_case(gn, solid=0.0, seed=11, gwc=None, acc=(1e-5, -2e-5, 3e-5)):
    """Two-component Shan-Chen case with random densities (and a random mask, solid sites initialised as
    initEqDistWall does): (params, populations, mask)."""
    from dlbc_b200 import init as I
    from dlbc_b200.connectivity import d3q19
    rng = np.random.default_rng(seed)
    mask = (rng.random(gn) < solid).astype(np.int32) if solid > 0 else None
    pops = []
    for rho in (0.6, 0.12):
        a = (rho + 0.05 * rng.random(gn))[..., None] * I.eq_dist_unity(d3q19, I.BDIST2)
        pops.append(I.init_eq_dist_wall(a, mask) if mask is not None else a)
    params = dict(gn=gn, components=2, global_acc=list(acc), **SC)
    if gwc is not None:
        params["gwc"] = list(gwc)
    return params, pops, mask
