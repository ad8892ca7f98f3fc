"""ctypes wrapper around the CPU oracle (oracle/dlbc_oracle.c).

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package dlbc_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libdlbc_oracle.so")
MAXC = 8

PSI = {"Linear": 0, "Exponential": 1}
EQ = {"SecondOrder": 0, "ThirdOrder": 1, "BDist2": 2}


class _Config(C.Structure):
    _fields_ = [
        ("d", C.c_int), ("q", C.c_int), ("n", C.c_int64 * 3), ("ncomp", C.c_int),
        ("tau", C.c_double * MAXC), ("gcc", C.c_double * (MAXC * MAXC)), ("gwc", C.c_double * MAXC),
        ("global_acc", C.c_double * 3), ("psi_form", C.c_int), ("eqdist_form", C.c_int),
        ("enable_shan_chen", C.c_int), ("nthreads", C.c_int),
    ]


def build(force=False):
    """Compile the oracle with oracle/Makefile (gcc).  Building the checker is not using it."""
    srcs = [os.path.join(_HERE, f) for f in ("dlbc_oracle.c", "dlbc_oracle_conn.inl", "dlbc_oracle.h", "Makefile")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int32)
        L.oracle_create.restype = C.c_void_p
        L.oracle_create.argtypes = [C.POINTER(_Config)]
        L.oracle_destroy.argtypes = [C.c_void_p]
        L.oracle_sites_with_halo.restype = C.c_int64
        L.oracle_sites_with_halo.argtypes = [C.c_void_p]
        L.oracle_set_populations.argtypes = [C.c_void_p, C.c_int, dp, C.c_int]
        L.oracle_get_populations.argtypes = [C.c_void_p, C.c_int, dp, C.c_int]
        L.oracle_set_mask.argtypes = [C.c_void_p, ip, C.c_int]
        L.oracle_get_mask.argtypes = [C.c_void_p, ip, C.c_int]
        for name in ("exchange_halo", "advect_fields", "reset_force", "add_shan_chen_force",
                     "distribute_force", "prepare_to_collide", "collide_fields"):
            getattr(L, "oracle_" + name).argtypes = [C.c_void_p]
        L.oracle_step.argtypes = [C.c_void_p, C.c_int]
        for name in ("density", "velocity", "force", "psi"):
            getattr(L, "oracle_" + name).argtypes = [C.c_void_p, C.c_int, dp]
        L.oracle_colour.argtypes = [C.c_void_p, C.c_int, C.c_int, dp]
        L.oracle_total_mass.restype = C.c_double
        L.oracle_total_mass.argtypes = [C.c_void_p, C.c_int]
        L.oracle_total_momentum.argtypes = [C.c_void_p, C.c_int, dp]
        L.oracle_site_density.restype = C.c_double
        L.oracle_site_density.argtypes = [C.c_int, dp]
        L.oracle_site_velocity.argtypes = [C.c_int, C.c_int, dp, C.c_double, dp]
        L.oracle_site_eqdist.argtypes = [C.c_int, C.c_int, C.c_int, dp, dp, C.c_double, dp]
        L.oracle_site_psi.restype = C.c_double
        L.oracle_site_psi.argtypes = [C.c_int, C.c_double]
        L.oracle_tables.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), dp, dp]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class OracleLattice:
    """One periodic lattice advanced by the CPU restatement of runTimeloop (main.d:174-200)."""

    def __init__(self, gn, components=1, tau=None, gcc=None, gwc=None, global_acc=None,
                 psi_form="Linear", eqdist_form="SecondOrder", enable_shan_chen=False, nthreads=0):
        gn = tuple(int(x) for x in gn)
        self.d = len(gn)
        self.q = {2: 9, 3: 19}[self.d]
        self.gn = gn
        self.ncomp = int(components)
        cfg = _Config()
        cfg.d, cfg.q, cfg.ncomp = self.d, self.q, self.ncomp
        for i in range(3):
            cfg.n[i] = gn[i] if i < self.d else 1
        tau = list(tau) if tau is not None else [1.0] * self.ncomp
        for i, t in enumerate(tau):
            cfg.tau[i] = t
        if gcc is not None:
            for i in range(self.ncomp):
                for j in range(self.ncomp):
                    cfg.gcc[i * MAXC + j] = gcc[i][j]
        if gwc is not None:
            for i, g in enumerate(gwc):
                cfg.gwc[i] = g
        if global_acc is not None:
            for i, g in enumerate(global_acc):
                cfg.global_acc[i] = g
        cfg.psi_form = PSI[psi_form] if isinstance(psi_form, str) else int(psi_form)
        cfg.eqdist_form = EQ[eqdist_form] if isinstance(eqdist_form, str) else int(eqdist_form)
        cfg.enable_shan_chen = int(bool(enable_shan_chen))
        cfg.nthreads = int(nthreads)
        self._l = lib()
        self._h = self._l.oracle_create(C.byref(cfg))
        if not self._h:
            raise ValueError("oracle_create rejected the configuration")

    def __del__(self):
        if getattr(self, "_h", None):
            self._l.oracle_destroy(self._h)
            self._h = None

    def _shape(self, halo, tail=()):
        return tuple(g + 2 * halo for g in self.gn) + tuple(tail)

    def set_populations(self, comp, aos, halo=0):
        a = np.ascontiguousarray(aos, dtype=np.float64)
        assert a.shape == self._shape(halo, (self.q,)), (a.shape, self._shape(halo, (self.q,)))
        self._l.oracle_set_populations(self._h, comp, _dp(a), halo)

    def get_populations(self, comp, halo=0):
        out = np.empty(self._shape(halo, (self.q,)))
        self._l.oracle_get_populations(self._h, comp, _dp(out), halo)
        return out

    def set_mask(self, mask, halo=0):
        m = np.ascontiguousarray(mask, dtype=np.int32)
        assert m.shape == self._shape(halo)
        self._l.oracle_set_mask(self._h, m.ctypes.data_as(C.POINTER(C.c_int32)), halo)

    def get_mask(self, halo=0):
        out = np.empty(self._shape(halo), dtype=np.int32)
        self._l.oracle_get_mask(self._h, out.ctypes.data_as(C.POINTER(C.c_int32)), halo)
        return out

    def exchange_halo(self): self._l.oracle_exchange_halo(self._h)
    def advect_fields(self): self._l.oracle_advect_fields(self._h)
    def reset_force(self): self._l.oracle_reset_force(self._h)
    def add_shan_chen_force(self): self._l.oracle_add_shan_chen_force(self._h)
    def distribute_force(self): self._l.oracle_distribute_force(self._h)
    def prepare_to_collide(self): self._l.oracle_prepare_to_collide(self._h)
    def collide_fields(self): self._l.oracle_collide_fields(self._h)

    def step(self, nsteps=1):
        self._l.oracle_step(self._h, int(nsteps))

    def density(self, comp):
        out = np.empty(self.gn)
        self._l.oracle_density(self._h, comp, _dp(out))
        return out

    def psi(self, comp):
        out = np.empty(self.gn)
        self._l.oracle_psi(self._h, comp, _dp(out))
        return out

    def colour(self, c1, c2):
        out = np.empty(self.gn)
        self._l.oracle_colour(self._h, c1, c2, _dp(out))
        return out

    def velocity(self, comp):
        out = np.empty(self.gn + (self.d,))
        self._l.oracle_velocity(self._h, comp, _dp(out))
        return out

    def force(self, comp):
        out = np.empty(self.gn + (self.d,))
        self._l.oracle_force(self._h, comp, _dp(out))
        return out

    def total_mass(self, comp):
        return float(self._l.oracle_total_mass(self._h, comp))

    def total_momentum(self, comp):
        out = np.empty(3)
        self._l.oracle_total_momentum(self._h, comp, _dp(out))
        return out[:self.d]


# ---- per-site functions ---------------------------------------------------------------------
def site_density(pop):
    p = np.ascontiguousarray(pop, dtype=np.float64)
    return float(lib().oracle_site_density(len(p), _dp(p)))


def site_velocity(pop, rho=None, d=3):
    p = np.ascontiguousarray(pop, dtype=np.float64)
    rho = site_density(p) if rho is None else rho
    out = np.zeros(3)
    lib().oracle_site_velocity(d, len(p), _dp(p), rho, _dp(out))
    return out[:d]


def site_eqdist(form, pop, dv, rho0=None, d=3):
    p = np.ascontiguousarray(pop, dtype=np.float64)
    v = np.zeros(3)
    v[:d] = dv
    rho0 = site_density(p) if rho0 is None else rho0
    out = np.empty(len(p))
    lib().oracle_site_eqdist(d, len(p), EQ[form] if isinstance(form, str) else form, _dp(p), _dp(v), rho0, _dp(out))
    return out


def site_psi(form, rho):
    return float(lib().oracle_site_psi(PSI[form] if isinstance(form, str) else form, rho))


def tables(d, q):
    v = np.empty((q, d), dtype=np.int64)
    b = np.empty(q, dtype=np.int64)
    w = np.empty(q)
    css = C.c_double()
    lib().oracle_tables(d, q, v.ctypes.data_as(C.POINTER(C.c_int64)), b.ctypes.data_as(C.POINTER(C.c_int64)),
                        _dp(w), C.byref(css))
    return v, b, w, css.value
