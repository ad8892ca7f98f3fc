"""Laplace-law measurement of the surface tension (reference src/dlbc/lb/laplace.d:55-222, 3-D branch).

dumpLaplace needs four numbers from the lattice: the volume-averaged densities in a small cube at the centre of
the (centred) droplet and in one next to the origin, and the total masses.  `Lattice.volume_averaged_density`
and `Lattice.total_mass` compute them on the device; the rest is scalar arithmetic, restated here in the
reference's order.  `report_line` formats a result as writeToFile does (laplace.d:277-296), so the output can be
compared with the reference's ``laplace-*.asc`` files.
"""
from __future__ import annotations

import math


def pressure(density, gcc, psi_form="Exponential", css=1.0 / 3.0):
    """density.d:306-326 for a list of component densities."""
    psi = (lambda r: r) if psi_form == "Linear" else (lambda r: 1.0 - math.exp(-r))
    p = 0.0
    for i, d1 in enumerate(density):
        p += d1
        for j, d2 in enumerate(density):
            p += 0.5 * gcc[i][j] * psi(d1) * psi(d2)
    return p * css


def measure(L, gn, gcc, psi_form="Exponential", volume_average=1, previous_sigma=float("nan")):
    """One evaluation of dumpLaplace on lattice `L` (two components, 3-D).  Returns a dict with sigma, rel,
    measuredR, inPres, outPres (deformation D is 2-D only in the reference: NaN)."""
    offset_in = [int(n * 0.5) for n in gn]
    offset_out = [volume_average] * 3
    in_den = L.volume_averaged_density(offset_in, volume_average)
    out_den = L.volume_averaged_density(offset_out, volume_average)
    in_pres = pressure(list(in_den), gcc, psi_form)
    out_pres = pressure(list(out_den), gcc, psi_form)
    g_mass = [L.total_mass(c) for c in range(2)]
    gsize = gn[0] * gn[1] * gn[2]
    eff_mass = g_mass[0] - gsize * out_den[0]
    with_err = eff_mass / (4.0 / 3.0 * math.pi * (in_den[0] - out_den[0]))
    measured_r = math.copysign(abs(with_err) ** (1.0 / 3.0), with_err) if with_err == with_err else float("nan")
    sigma = measured_r * (in_pres - out_pres) / 2.0
    try:
        rel = sigma / previous_sigma - 1.0
    except ZeroDivisionError:
        rel = float("inf") if sigma > 0 else float("nan")
    return dict(sigma=sigma, rel=rel, measuredR=measured_r, inPres=in_pres, outPres=out_pres, D=float("nan"),
                inDen=list(in_den), outDen=list(out_den), mass=g_mass)


def report_line(gcc01, init_r, t, m):
    """writeToFile's data line (laplace.d:294): "%e %e %8d %+e %+e %e %e %e %e"."""
    return "%e %e %8d %+e %+e %e %e %e %e" % (gcc01, init_r, t, m["sigma"], m["rel"], m["measuredR"], m["inPres"],
                                                m["outPres"], m["D"])
