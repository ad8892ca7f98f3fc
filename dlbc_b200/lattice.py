"""Host-side mirror of the reference's lattice interface for the timestep path.

``Lattice`` plays the role of ``Lattice!(conn)`` (src/dlbc/lattice.d:95-154) for a host program:
it is constructed from the same parameters the reference reads from its parameter file
(lattice.gn, lb.components, lb.tau, lb.force.*, lb.eqdist.eqDistForm), receives the arrays that
initFluids / initMask / readCheckpoint produce, and `step()` stands for the statements of
runTimeloop (src/main.d:179-195).  Everything goes through the C ABI of libdlbc_cuda.so; this
module holds no arithmetic of its own.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib as _lib

Q = 19


def slab_range(gnz: int, rank: int, nranks: int):
    """z-slab owned by `rank`: the reference's M.nc = [1,1,nranks], M.c = [0,0,rank]
    (parallel.d:190-237, lattice.d:183-187)."""
    if gnz % nranks:
        raise ValueError(f"gn[2] = {gnz} is not divisible by {nranks} ranks (lattice.d:233-241)")
    n = gnz // nranks
    return rank * n, (rank + 1) * n


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class Lattice:
    def __init__(self, gn, components=1, tau=None, gcc=None, gwc=None, global_acc=None, psi_form="Linear",
                 eqdist_form="SecondOrder", enable_shan_chen=False, arithmetic="fast", kernel="auto",
                 device=0, rank=0, nranks=1, nccl_id=None):
        self._l = _lib.load()
        self._h = None
        gn = tuple(int(x) for x in gn)
        if len(gn) not in (2, 3):
            raise ValueError("the CUDA path is built for D3Q19 (gn with 3 entries) and D2Q9 (2 entries)")
        self.gn = gn
        self.d = len(gn)
        self.q = Q if self.d == 3 else 9  # Lattice!(d3q19) / Lattice!(d2q9), connectivity.d:21-32
        self.ncomp = int(components)
        self.rank, self.nranks = int(rank), int(nranks)
        if self.d == 3:
            z0, z1 = slab_range(gn[2], self.rank, self.nranks)
            self.z_range = (z0, z1)
            self.n = (gn[0], gn[1], z1 - z0)  # local physical extents
        else:
            if self.nranks != 1:
                raise ValueError("a 2-D lattice runs on one rank")
            self.z_range = (0, 1)
            self.n = gn
        cfg = _lib.Config()
        cfg.abi_version = _lib.ABI_VERSION
        cfg.d, cfg.q = self.d, self.q
        for i in range(self.d):
            cfg.gn[i] = gn[i]
        cfg.ncomp = self.ncomp
        if self.ncomp < 1 or self.ncomp > _lib.MAXC:
            raise ValueError(f"components must be in 1..{_lib.MAXC}")
        tau = [1.0] * self.ncomp if tau is None else list(tau)
        if len(tau) != self.ncomp:
            raise ValueError("lb.tau needs one entry per component")
        for i, t in enumerate(tau):
            cfg.tau[i] = t
        if gcc is not None:
            # prepareForce (force.d:67-107): gcc must be components x components
            if len(gcc) != self.ncomp or any(len(r) != self.ncomp for r in gcc):
                raise ValueError("lb.force.gcc must be a components x components matrix (force.d:80-91)")
            for i in range(self.ncomp):
                for j in range(self.ncomp):
                    cfg.gcc[i * _lib.MAXC + j] = gcc[i][j]
        if gwc is not None:
            if len(gwc) != self.ncomp:
                raise ValueError("lb.force.gwc needs one entry per component (force.d:98-103)")
            for i, g in enumerate(gwc):
                cfg.gwc[i] = g
        if global_acc is not None:
            for i, g in enumerate(global_acc):
                cfg.global_acc[i] = g
        cfg.psi_form = _lib.PSI[psi_form] if isinstance(psi_form, str) else int(psi_form)
        cfg.eqdist_form = _lib.EQ[eqdist_form] if isinstance(eqdist_form, str) else int(eqdist_form)
        cfg.enable_shan_chen = int(bool(enable_shan_chen))
        cfg.arithmetic = _lib.ARITH[arithmetic] if isinstance(arithmetic, str) else int(arithmetic)
        cfg.kernel_variant = _lib.KERNEL[kernel] if isinstance(kernel, str) else int(kernel)
        cfg.device = int(device)
        cfg.rank, cfg.nranks = self.rank, self.nranks
        if self.nranks > 1:
            if nccl_id is None or len(nccl_id) != _lib.NCCL_ID_BYTES:
                raise ValueError("nranks > 1 needs the 128-byte nccl_id from lib.nccl_unique_id() of rank 0")
            C.memmove(cfg.nccl_id, bytes(nccl_id), _lib.NCCL_ID_BYTES)
        h = C.c_void_p()
        rc = self._l.dlbc_cuda_create(C.byref(cfg), C.byref(h))
        if rc:
            raise _lib.DlbcCudaError(f"dlbc_cuda_create failed ({rc}): {self._l.dlbc_cuda_last_error(None).decode()}")
        self._h = h

    # ---- plumbing -----------------------------------------------------------------------------
    def _ck(self, rc):
        if rc:
            raise _lib.DlbcCudaError(f"error {rc}: {self._l.dlbc_cuda_last_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None):
            self._l.dlbc_cuda_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _shape(self, halo, tail=()):
        return tuple(x + 2 * halo for x in self.n) + tuple(tail)

    # ---- state --------------------------------------------------------------------------------
    def set_populations(self, comp, aos, halo=0):
        """L.fluids[comp] <- aos ([x][y][z][19] incl. `halo` layers: Field.arr when halo = 2)."""
        a = np.ascontiguousarray(aos, dtype=np.float64)
        if a.shape != self._shape(halo, (self.q,)):
            raise ValueError(f"populations shape {a.shape} != {self._shape(halo, (self.q,))}")
        self._ck(self._l.dlbc_cuda_upload_populations(self._h, comp, _dp(a), halo))

    def get_populations(self, comp, halo=0):
        out = np.zeros(self._shape(halo, (self.q,)))
        self._ck(self._l.dlbc_cuda_download_populations(self._h, comp, _dp(out), halo))
        return out

    def set_populations_xrange(self, comp, layers, x_begin, halo=0):
        a = np.ascontiguousarray(layers, dtype=np.float64)
        want = (a.shape[0],) + self._shape(halo, (Q,))[1:]
        if a.shape != want:
            raise ValueError(f"layer block shape {a.shape} != {want}")
        self._ck(self._l.dlbc_cuda_upload_populations_xrange(self._h, comp, _dp(a), halo, x_begin, a.shape[0]))

    def get_populations_xrange(self, comp, x_begin, x_count, halo=0):
        out = np.zeros((x_count,) + self._shape(halo, (Q,))[1:])
        self._ck(self._l.dlbc_cuda_download_populations_xrange(self._h, comp, _dp(out), halo, x_begin, x_count))
        return out

    def set_mask(self, mask, halo=0):
        m = np.ascontiguousarray(mask, dtype=np.int32)
        if m.shape != self._shape(halo):
            raise ValueError(f"mask shape {m.shape} != {self._shape(halo)}")
        self._ck(self._l.dlbc_cuda_upload_mask(self._h, m.ctypes.data_as(C.POINTER(C.c_int32)), halo))

    def get_mask(self, halo=0):
        out = np.zeros(self._shape(halo), dtype=np.int32)
        self._ck(self._l.dlbc_cuda_download_mask(self._h, out.ctypes.data_as(C.POINTER(C.c_int32)), halo))
        return out

    # ---- time loop ----------------------------------------------------------------------------
    def step(self, nsteps=1, sync=True):
        self._ck(self._l.dlbc_cuda_step(self._h, int(nsteps)))
        if sync:
            self.sync()

    def sync(self):
        self._ck(self._l.dlbc_cuda_sync(self._h))

    @property
    def timestep(self):
        return int(self._l.dlbc_cuda_timestep(self._h))

    # ---- outputs (dumpLBData, lb/io.d:56-101) ---------------------------------------------------
    def density(self, comp):
        out = np.empty(self.n)
        self._ck(self._l.dlbc_cuda_download_density(self._h, comp, _dp(out), 0))
        return out

    def colour(self, c1, c2):
        out = np.empty(self.n)
        self._ck(self._l.dlbc_cuda_download_colour(self._h, c1, c2, _dp(out), 0))
        return out

    def velocity(self, comp):
        out = np.empty(self.n + (self.d,))
        self._ck(self._l.dlbc_cuda_download_velocity(self._h, comp, _dp(out), 0))
        return out

    def force(self, comp):
        out = np.empty(self.n + (self.d,))
        self._ck(self._l.dlbc_cuda_download_force(self._h, comp, _dp(out), 0))
        return out

    def pressure(self):
        """density.d:306-326 on every site."""
        out = np.empty(self.n)
        self._ck(self._l.dlbc_cuda_download_pressure(self._h, _dp(out), 0))
        return out

    def corrected_velocity(self, comp):
        """velocity.d:177-185: raw velocity + tau * F / rho (force of the last step), zero on solid sites."""
        out = np.empty(self.n + (3,))
        self._ck(self._l.dlbc_cuda_download_corrected_velocity(self._h, comp, _dp(out), 0))
        return out

    def volume_averaged_density(self, offset, half_width=1):
        """laplace.d:236-272: mean density of every component over the fluid sites of the cube offset +- half_width."""
        off = (C.c_int64 * 3)(*[int(v) for v in offset])
        out = np.empty(self.ncomp)
        self._ck(self._l.dlbc_cuda_volume_averaged_density(self._h, off, int(half_width), _dp(out)))
        return out

    # ---- geometry / initial conditions generated on the device (fields/init.d, lb/mask.d) ----------
    def init_mask(self, kind="None", axis=0, wall_offset=0, thickness=1, dimension_has_wall=None):
        hw = None
        if dimension_has_wall is not None:
            flat = [int(bool(v)) for row in dimension_has_wall for v in row]
            hw = (C.c_int32 * 6)(*(flat + [0] * 6)[:6])
        self._ck(self._l.dlbc_cuda_init_mask(self._h, _lib.MASK_KIND[kind], int(axis), int(wall_offset), int(thickness), hw))

    def init_populations(self, comp, kind, density, perturb=0.0, seed=0, skip=0):
        self._ck(self._l.dlbc_cuda_init_populations(self._h, int(comp), _lib.INIT_KIND[kind], float(density), float(perturb),
                                                    int(seed) & 0xFFFFFFFF, int(skip)))

    def init_spheres(self, comp, density1, density2, radius, thickness, offset=(0, 0, 0), separation=None, cylinder_axis=None,
                     frac=False):
        """initEqDistSphere / TwoSpheres (separation given) / Cylinder (cylinder_axis given) and their *Frac variants
        (fields/init.d:84-259); the scalar arithmetic on the parameters is dlbc_b200.init's."""
        gn = self.gn
        if frac:
            small = min(g for i, g in enumerate(gn) if i != cylinder_axis)
            radius = radius * small
            offset = [offset[i] * gn[i] for i in range(3)]
            if separation is not None:
                separation = [separation[i] * gn[i] for i in range(3)]
        if separation is None:
            centres = [[1 * gn[i] * 0.5 + offset[i] for i in range(3)]]
        else:
            centres = [[1 * gn[i] * 0.5 + offset[i] - 0.5 * separation[i] for i in range(3)],
                       [1 * gn[i] * 0.5 + offset[i] + 0.5 * separation[i] for i in range(3)]]
        flat = np.array(centres, dtype=np.float64).ravel()
        self._ck(self._l.dlbc_cuda_init_spheres(self._h, int(comp), len(centres), _dp(flat),
                                                -1 if cylinder_axis is None else int(cylinder_axis), float(radius),
                                                float(thickness), float(density1), float(density2)))

    def init_lamellae(self, comp, values, widths, axis, thickness, frac=False):
        """initEqDistLamellae(Frac) (fields/init.d:261-313): the 1-D profile is evaluated on the host."""
        from . import init as I
        if frac:
            widths = [w * self.gn[axis] for w in widths]
        line = I.Block((self.gn[axis],) * 3, halo=0)
        prof = np.ascontiguousarray(I.lamellae_values(line, values, widths, 0, thickness)[:, 0, 0], dtype=np.float64)
        self._ck(self._l.dlbc_cuda_init_axis_profile(self._h, int(comp), int(axis), _dp(prof)))

    def init_wall(self):
        self._ck(self._l.dlbc_cuda_init_wall(self._h))

    def total_mass(self, comp):
        v = C.c_double()
        self._ck(self._l.dlbc_cuda_total_mass(self._h, comp, C.byref(v)))
        return v.value

    def total_momentum(self, comp):
        out = np.empty(3)
        self._ck(self._l.dlbc_cuda_total_momentum(self._h, comp, _dp(out)))
        return out

    # ---- measurement --------------------------------------------------------------------------
    def timed_steps(self, nsteps):
        """Run nsteps between two CUDA events on the compute stream; returns milliseconds."""
        ms = C.c_double()
        self._ck(self._l.dlbc_cuda_timed_steps(self._h, int(nsteps), C.byref(ms)))
        return ms.value

    def kernel_times(self):
        """One step with CUDA events around every kernel: [(name, ms), ...]."""
        cap = 16
        names = (C.c_char_p * cap)()
        ms = (C.c_double * cap)()
        n = C.c_int()
        self._ck(self._l.dlbc_cuda_kernel_times(self._h, cap, names, ms, C.byref(n)))
        return [(names[i].decode(), ms[i]) for i in range(n.value)]

    def set_option(self, name, value):
        """Tuning knobs of the single-launch schedules ("wave_tyb", "wave_la"); results do not depend on them."""
        self._ck(self._l.dlbc_cuda_set_option(self._h, name.encode(), int(value)))

    @property
    def kernel_variant(self):
        """Name of the step kernel variant the next step runs (what "auto" resolved to)."""
        v = C.c_int()
        self._l.dlbc_cuda_kernel_variant(self._h, C.byref(v))
        return {n: k for k, n in _lib.KERNEL.items()}[v.value]

    @property
    def launch_count(self):
        v = C.c_int64()
        self._l.dlbc_cuda_launch_count(self._h, C.byref(v))
        return v.value

    @property
    def device_bytes(self):
        v = C.c_int64()
        self._l.dlbc_cuda_device_bytes(self._h, C.byref(v))
        return v.value
