"""Host-side initial conditions for populations and masks (numpy).

The north-star keeps mask/geometry setup and initial conditions on the host; the CUDA
library only receives the finished arrays.  These functions restate the reference's
initialisers so the same ``.in`` files produce bit-identical inputs:

* src/dlbc/fields/init.d:26-82      initConst, initConstRandom, initEqDist, initEqDistRandom,
                                    initEqDistPerturb, initEqDistPerturbFrac
* src/dlbc/fields/init.d:84-197     spheres / two spheres (+Frac variants)
* src/dlbc/fields/init.d:199-259    cylinders (+Frac)
* src/dlbc/fields/init.d:261-313    lamellae (+Frac), :315-374 helpers
* src/dlbc/fields/init.d:377-467    initBox, initTube, initPlates, initWalls
* src/dlbc/fields/init.d:469-477    initEqDistWall
* src/dlbc/lb/eqdist.d:310-315      eqDistUnity
* src/dlbc/lb/init.d:109-125        per-fluid reseeding

Array convention (same as the reference's ``Field.arr``): local block *including* a halo of
``h`` sites on every side, C order, populations AoS ``[x][y][z][q]``.  ``c``/``nc`` are the
Cartesian coordinates / grid of the emulated MPI rank (``M.c``, ``M.nc``); a single-process
run is ``c = 0, nc = 1``.  Arithmetic is written in the reference's operation order so the
results are bit-identical (numpy float64 ops are IEEE, sqrt is correctly rounded).
"""
from __future__ import annotations

import numpy as np

from .connectivity import Connectivity

SECOND_ORDER, THIRD_ORDER, BDIST2 = 0, 1, 2
EQDIST_FORMS = {"SecondOrder": SECOND_ORDER, "ThirdOrder": THIRD_ORDER, "BDist2": BDIST2}


def eq_dist_unity(conn: Connectivity, form: int) -> np.ndarray:
    """eqDistUnity (eqdist.d:310-315): eqDist of pop=[1,0,...] with dv=0."""
    if form in (SECOND_ORDER, THIRD_ORDER):
        # rho0 = 1, v = 0: rho0 * w * (1 + 0 + 0 - 0 [+ 0 - 0])
        return (1.0 * conn.weights) * (1.0 + 0.0 + 0.0 - 0.0)
    if form == BDIST2:
        if conn.name == "d3q19":
            aN1, aN0 = 1.0 / 36.0, 1.0 / 3.0
            axis, diag = range(1, 7), range(7, 19)
        elif conn.name == "d2q9":
            aN1, aN0 = 1.0 / 24.0, 1.0 / 2.0
            axis, diag = range(1, 5), range(5, 9)
        else:
            raise ValueError("BDist2 only for d3q19/d2q9")
        out = np.empty(conn.q)
        cst = 1.0 - 0.0
        const1 = aN1 * cst
        out[0] = aN0 * cst
        for q in axis:
            out[q] = 2.0 * (const1 + 0.0)
        for q in diag:
            out[q] = const1 + 0.0
        return out * 1.0
    raise ValueError(f"unknown eqDistForm {form}")


def _coords(shape_h, axis):
    """Index p[axis] broadcast over the spatial block (as int64)."""
    d = len(shape_h)
    idx = np.arange(shape_h[axis], dtype=np.int64)
    return idx.reshape([-1 if i == axis else 1 for i in range(d)])


def _gn(shape_h, axis, c, n, halo):
    # gn[i] = p[i] + M.c[i] * field.n[i] - to!double(field.haloSize)
    return (_coords(shape_h, axis) + int(c[axis]) * int(n[axis])).astype(np.float64) - float(halo)


def symmetric_linear_transition(pos, width, negval, posval):
    """fields/init.d:346-366."""
    pos = np.asarray(pos, dtype=np.float64)
    if width == 0.0:
        return np.where(pos < 0, np.float64(negval), np.float64(posval))
    with np.errstate(all="ignore"):
        mid = negval * (0.5 - pos / width) + posval * (0.5 + pos / width)
    return np.where(pos < -0.5 * width, np.float64(negval),
                    np.where(pos > 0.5 * width, np.float64(posval), mid))


class Block:
    """Geometry of one (emulated) rank's block."""

    def __init__(self, gn, halo=2, nc=None, c=None):
        self.gn = tuple(int(g) for g in gn)
        self.d = len(self.gn)
        self.nc = tuple(nc) if nc is not None else (1,) * self.d
        self.c = tuple(c) if c is not None else (0,) * self.d
        for g, k in zip(self.gn, self.nc):
            if g % k:
                raise ValueError("gn must be divisible by nc (lattice.d:233-241)")
        self.n = tuple(g // k for g, k in zip(self.gn, self.nc))
        self.halo = int(halo)
        self.shape_h = tuple(x + 2 * self.halo for x in self.n)

    def interior(self):
        h = self.halo
        return tuple(slice(h, h + x) for x in self.n)


# ---- population initialisers -------------------------------------------------------------

def init_const(blk, q, fill):
    return np.full(blk.shape_h + (q,), np.float64(fill))


def init_const_random(blk, q, fill, lrng):
    nsite = int(np.prod(blk.shape_h))
    return (np.float64(fill) * lrng.uniform(0.0, 2.0, nsite * q)).reshape(blk.shape_h + (q,))


def init_eq_dist(blk, conn, form, density):
    eqpop = np.float64(density) * eq_dist_unity(conn, form)
    return np.broadcast_to(eqpop, blk.shape_h + (conn.q,)).copy()


def init_eq_dist_random(blk, conn, form, density, lrng):
    u = lrng.uniform(0.0, 2.0, int(np.prod(blk.shape_h))).reshape(blk.shape_h)
    return (np.float64(density) * u)[..., None] * eq_dist_unity(conn, form)


def init_eq_dist_perturb(blk, conn, form, density, perturb, lrng):
    u = lrng.uniform(-perturb, perturb, int(np.prod(blk.shape_h))).reshape(blk.shape_h)
    return (np.float64(density) + u)[..., None] * eq_dist_unity(conn, form)


def init_eq_dist_perturb_frac(blk, conn, form, density, perturb, lrng):
    u = lrng.uniform(-perturb, perturb, int(np.prod(blk.shape_h))).reshape(blk.shape_h)
    return (np.float64(density) * (1.0 + u))[..., None] * eq_dist_unity(conn, form)


def _dist_to(blk, centre_shift, skip_axis=None):
    """sqrt(offset . offset) with offset[i] = gn[i] - (nc*n*0.5 + shift[i]) + 0.5."""
    acc = np.float64(0.0)
    for i in range(blk.d):
        if i == skip_axis:
            off = np.float64(0.0)
        else:
            g = _gn(blk.shape_h, i, blk.c, blk.n, blk.halo)
            off = g - np.float64(centre_shift[i]) + 0.5
        acc = acc + off * off
    return np.sqrt(np.broadcast_to(acc, blk.shape_h))


def init_eq_dist_sphere(blk, conn, form, d1, d2, radius, offset, thickness):
    shift = [blk.nc[i] * blk.n[i] * 0.5 + offset[i] for i in range(blk.d)]
    relpos = _dist_to(blk, shift) - radius
    val = symmetric_linear_transition(relpos, thickness, d1, d2)
    return val[..., None] * eq_dist_unity(conn, form)


def _small_size(gn, skip=None):
    return min(g for i, g in enumerate(gn) if i != skip)


def init_eq_dist_sphere_frac(blk, conn, form, d1, d2, radius_frac, offset_frac, thickness):
    radius = radius_frac * _small_size(blk.gn)
    offset = [offset_frac[i] * blk.gn[i] for i in range(blk.d)]
    return init_eq_dist_sphere(blk, conn, form, d1, d2, radius, offset, thickness)


def init_eq_dist_two_spheres(blk, conn, form, d1, d2, radius, offset, thickness, separation):
    s1 = [blk.nc[i] * blk.n[i] * 0.5 + offset[i] - 0.5 * separation[i] for i in range(blk.d)]
    s2 = [blk.nc[i] * blk.n[i] * 0.5 + offset[i] + 0.5 * separation[i] for i in range(blk.d)]
    r1 = _dist_to(blk, s1) - radius
    r2 = _dist_to(blk, s2) - radius
    relpos = np.minimum(r1, r2)
    val = symmetric_linear_transition(relpos, thickness, d1, d2)
    return val[..., None] * eq_dist_unity(conn, form)


def init_eq_dist_two_spheres_frac(blk, conn, form, d1, d2, radius_frac, offset_frac, thickness, sep_frac):
    radius = radius_frac * _small_size(blk.gn)
    offset = [offset_frac[i] * blk.gn[i] for i in range(blk.d)]
    sep = [sep_frac[i] * blk.gn[i] for i in range(blk.d)]
    return init_eq_dist_two_spheres(blk, conn, form, d1, d2, radius, offset, thickness, sep)


def init_eq_dist_cylinder(blk, conn, form, d1, d2, axis, radius, offset, thickness):
    shift = [blk.nc[i] * blk.n[i] * 0.5 + offset[i] for i in range(blk.d)]
    relpos = _dist_to(blk, shift, skip_axis=axis) - radius
    val = symmetric_linear_transition(relpos, thickness, d1, d2)
    return val[..., None] * eq_dist_unity(conn, form)


def init_eq_dist_cylinder_frac(blk, conn, form, d1, d2, axis, radius_frac, offset_frac, thickness):
    radius = radius_frac * _small_size(blk.gn, skip=axis)
    offset = [offset_frac[i] * blk.gn[i] for i in range(blk.d)]
    return init_eq_dist_cylinder(blk, conn, form, d1, d2, axis, radius, offset, thickness)


def _calculate_interfaces(widths):
    itf = [0.0] + [float(w) for w in widths]
    for i in range(1, len(itf)):
        itf[i] += itf[i - 1]
    return np.array(itf) - 0.5


def lamellae_values(blk, values, widths, axis, thickness):
    """initLamellae (fields/init.d:288-313): scalar value per site."""
    values = [float(v) for v in values]
    assert len(values) == len(widths)
    itf = _calculate_interfaces(widths)
    gp = _gn(blk.shape_h, axis, blk.c, blk.n, blk.halo).ravel()
    out = np.empty_like(gp)
    for k, g in enumerate(gp):
        diffs = np.abs(itf - g)
        ci = int(np.argmin(diffs))  # first minimum == strict '<' scan
        md = g - itf[ci]
        if ci == 0:
            neg, pos = values[-1], values[ci]
        elif ci == len(values):
            neg, pos = values[ci - 1], values[0]
        else:
            neg, pos = values[ci - 1], values[ci]
        out[k] = symmetric_linear_transition(md, thickness, neg, pos)
    shape = [1] * blk.d
    shape[axis] = -1
    return np.broadcast_to(out.reshape(shape), blk.shape_h)


def init_eq_dist_lamellae(blk, conn, form, values, widths, axis, thickness):
    if axis >= blk.d:
        raise ValueError("lb.init.initAxis out of range for EqDistLamellae (lb/init.d:333-335)")
    val = lamellae_values(blk, values, widths, axis, thickness)
    return val[..., None] * eq_dist_unity(conn, form)


def init_eq_dist_lamellae_frac(blk, conn, form, values, widths_frac, axis, thickness):
    widths = [w * blk.gn[axis] for w in widths_frac]
    return init_eq_dist_lamellae(blk, conn, form, values, widths, axis, thickness)


def init_eq_dist_wall(pop, mask):
    """initEqDistWall (fields/init.d:469-477): solid sites get [1, 0, ..., 0]."""
    solid = mask == 1
    pop[solid] = 0.0
    pop[solid, 0] = 1.0
    return pop


# ---- mask initialisers -------------------------------------------------------------------

MASK_NONE, MASK_SOLID = 0, 1


def _gni(blk, axis):
    return _coords(blk.shape_h, axis) + blk.c[axis] * blk.n[axis] - blk.halo


def mask_none(blk):
    return np.zeros(blk.shape_h, dtype=np.int32)


def mask_box(blk):
    m = np.zeros(blk.shape_h, dtype=np.int32)
    for i in range(blk.d):
        g = _gni(blk, i)
        m |= ((g == 0) | (g == blk.n[i] * blk.nc[i] - 1)).astype(np.int32)
    return m


def mask_tube(blk, axis):
    m = np.zeros(blk.shape_h, dtype=np.int32)
    for i in range(blk.d):
        if i == axis:
            continue
        g = _gni(blk, i)
        m |= ((g == 0) | (g == blk.n[i] * blk.nc[i] - 1)).astype(np.int32)
    return m


def mask_plates(blk, axis, wall_offset=0):
    if axis >= blk.d:
        raise ValueError("lb.mask.initAxis out of range for MaskInit.Plates (mask.d:161-163)")
    g = _gni(blk, axis)
    w = (g == wall_offset) | (g == blk.n[axis] * blk.nc[axis] - 1 - wall_offset)
    return np.broadcast_to(w, blk.shape_h).astype(np.int32)


def mask_walls(blk, dimension_has_wall, thickness):
    m = np.zeros(blk.shape_h, dtype=np.int32)
    for i in range(blk.d):
        g = _gni(blk, i)
        if dimension_has_wall[i][0]:
            m |= (g < thickness).astype(np.int32)
        if dimension_has_wall[i][1]:
            m |= (g >= blk.n[i] * blk.nc[i] - thickness).astype(np.int32)
    return m


# ---- halo exchange on the host (periodic, single rank) -------------------------------------

def exchange_halo_periodic(arr, d, halo):
    """exchangeHalo for nc = 1 in every dimension (fields/parallel.d:58-103): x, then y, then z;
    each phase copies the ``halo`` outermost physical layers into the opposite halo layers,
    including halos already filled by earlier phases."""
    h = halo
    for ax in range(d):
        n = arr.shape[ax] - 2 * h
        lo = [slice(None)] * arr.ndim
        hi = [slice(None)] * arr.ndim
        src_lo = [slice(None)] * arr.ndim
        src_hi = [slice(None)] * arr.ndim
        lo[ax] = slice(0, h)
        hi[ax] = slice(h + n, 2 * h + n)
        src_hi[ax] = slice(n, n + h)      # upper physical layers -> lower halo
        src_lo[ax] = slice(h, 2 * h)      # lower physical layers -> upper halo
        arr[tuple(lo)] = arr[tuple(src_hi)]
        arr[tuple(hi)] = arr[tuple(src_lo)]
    return arr
