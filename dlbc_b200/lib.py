"""ctypes binding of libdlbc_cuda.so (the C ABI in include/dlbc_cuda.h).

This is the same binding a D host would write as an ``extern(C)`` block: plain
pointers and sizes, int return codes checked after every call.  There is no CPU fallback: if the
shared library has not been built, or no CUDA device is present, the calls fail loudly.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("DLBC_CUDA_LIB") or os.path.join(HERE, "libdlbc_cuda.so")  # override: A/B runs of two builds
MAXC = 4
NCCL_ID_BYTES = 128
ABI_VERSION = 3

PSI = {"Linear": 0, "Exponential": 1}
EQ = {"SecondOrder": 0, "ThirdOrder": 1, "BDist2": 2}
ARITH = {"fast": 0, "exact": 1}
KERNEL = {"auto": 0, "two_pass": 1, "wave": 2, "tma": 3}

# every symbol include/dlbc_cuda.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "dlbc_cuda_abi_version", "dlbc_cuda_device_count", "dlbc_cuda_nccl_unique_id", "dlbc_cuda_create",
    "dlbc_cuda_destroy", "dlbc_cuda_last_error", "dlbc_cuda_host_alloc", "dlbc_cuda_host_free",
    "dlbc_cuda_upload_populations", "dlbc_cuda_download_populations", "dlbc_cuda_upload_populations_xrange",
    "dlbc_cuda_download_populations_xrange", "dlbc_cuda_upload_mask", "dlbc_cuda_download_mask",
    "dlbc_cuda_step", "dlbc_cuda_sync", "dlbc_cuda_timestep", "dlbc_cuda_download_density",
    "dlbc_cuda_download_colour", "dlbc_cuda_download_velocity", "dlbc_cuda_download_force",
    "dlbc_cuda_total_mass", "dlbc_cuda_total_momentum", "dlbc_cuda_tables", "dlbc_cuda_launch_count",
    "dlbc_cuda_timed_steps", "dlbc_cuda_kernel_times", "dlbc_cuda_device_bytes", "dlbc_cuda_kernel_variant",
    "dlbc_cuda_set_option", "dlbc_cuda_download_pressure", "dlbc_cuda_download_corrected_velocity",
    "dlbc_cuda_volume_averaged_density", "dlbc_cuda_init_mask", "dlbc_cuda_init_populations", "dlbc_cuda_init_wall",
    "dlbc_cuda_init_spheres", "dlbc_cuda_init_axis_profile", "dlbc_cuda_site_function",
    "dlbc_cuda_mass_momentum_no_alloc",
]
MASK_KIND = {"None": 0, "Box": 1, "Tube": 2, "Plates": 3, "Walls": 4}
INIT_KIND = {"EqDist": 0, "EqDistRandom": 1, "EqDistPerturb": 2, "EqDistPerturbFrac": 3}


class Config(C.Structure):
    """struct dlbc_cuda_config"""
    _fields_ = [
        ("abi_version", C.c_int32), ("d", C.c_int32), ("q", C.c_int32), ("gn", C.c_int64 * 3),
        ("ncomp", C.c_int32), ("tau", C.c_double * MAXC), ("gcc", C.c_double * (MAXC * MAXC)),
        ("gwc", C.c_double * MAXC), ("global_acc", C.c_double * 3), ("psi_form", C.c_int32),
        ("eqdist_form", C.c_int32), ("enable_shan_chen", C.c_int32), ("arithmetic", C.c_int32),
        ("kernel_variant", C.c_int32), ("device", C.c_int32), ("rank", C.c_int32), ("nranks", C.c_int32),
        ("nccl_id", C.c_uint8 * NCCL_ID_BYTES),
    ]


class DlbcCudaError(RuntimeError):
    pass


_lib = None


def load():
    """dlopen libdlbc_cuda.so and declare the prototypes.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise DlbcCudaError(
            f"{SO_PATH} is missing: build it with `python -m dlbc_b200.build` (nvcc, sm_100a). "
            "dlbc_b200 has no CPU fallback.")
    L = C.CDLL(SO_PATH)
    vp, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32)
    i64p = C.POINTER(C.c_int64)
    L.dlbc_cuda_abi_version.restype = C.c_int
    L.dlbc_cuda_device_count.argtypes = [C.POINTER(C.c_int)]
    L.dlbc_cuda_nccl_unique_id.argtypes = [C.POINTER(C.c_uint8)]
    L.dlbc_cuda_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.dlbc_cuda_destroy.argtypes = [vp]
    L.dlbc_cuda_last_error.restype = C.c_char_p
    L.dlbc_cuda_last_error.argtypes = [vp]
    L.dlbc_cuda_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.dlbc_cuda_host_free.argtypes = [vp]
    L.dlbc_cuda_upload_populations.argtypes = [vp, C.c_int, dp, C.c_int]
    L.dlbc_cuda_download_populations.argtypes = [vp, C.c_int, dp, C.c_int]
    L.dlbc_cuda_upload_populations_xrange.argtypes = [vp, C.c_int, dp, C.c_int, C.c_int64, C.c_int64]
    L.dlbc_cuda_download_populations_xrange.argtypes = [vp, C.c_int, dp, C.c_int, C.c_int64, C.c_int64]
    L.dlbc_cuda_upload_mask.argtypes = [vp, ip, C.c_int]
    L.dlbc_cuda_download_mask.argtypes = [vp, ip, C.c_int]
    L.dlbc_cuda_step.argtypes = [vp, C.c_int]
    L.dlbc_cuda_sync.argtypes = [vp]
    L.dlbc_cuda_timestep.restype = C.c_int64
    L.dlbc_cuda_timestep.argtypes = [vp]
    L.dlbc_cuda_download_density.argtypes = [vp, C.c_int, dp, C.c_int]
    L.dlbc_cuda_download_colour.argtypes = [vp, C.c_int, C.c_int, dp, C.c_int]
    L.dlbc_cuda_download_velocity.argtypes = [vp, C.c_int, dp, C.c_int]
    L.dlbc_cuda_download_force.argtypes = [vp, C.c_int, dp, C.c_int]
    L.dlbc_cuda_total_mass.argtypes = [vp, C.c_int, dp]
    L.dlbc_cuda_total_momentum.argtypes = [vp, C.c_int, dp]
    L.dlbc_cuda_tables.argtypes = [i64p, i64p, dp, dp]
    L.dlbc_cuda_launch_count.argtypes = [vp, i64p]
    L.dlbc_cuda_timed_steps.argtypes = [vp, C.c_int, dp]
    L.dlbc_cuda_kernel_times.argtypes = [vp, C.c_int, C.POINTER(C.c_char_p), dp, C.POINTER(C.c_int)]
    L.dlbc_cuda_device_bytes.argtypes = [vp, i64p]
    L.dlbc_cuda_kernel_variant.argtypes = [vp, C.POINTER(C.c_int)]
    L.dlbc_cuda_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    L.dlbc_cuda_download_pressure.argtypes = [vp, dp, C.c_int]
    L.dlbc_cuda_download_corrected_velocity.argtypes = [vp, C.c_int, dp, C.c_int]
    L.dlbc_cuda_volume_averaged_density.argtypes = [vp, i64p, C.c_int, dp]
    L.dlbc_cuda_init_mask.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, ip]
    L.dlbc_cuda_init_populations.argtypes = [vp, C.c_int, C.c_int, C.c_double, C.c_double, C.c_uint32, C.c_int64]
    L.dlbc_cuda_init_wall.argtypes = [vp]
    L.dlbc_cuda_init_spheres.argtypes = [vp, C.c_int, C.c_int, dp, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double]
    L.dlbc_cuda_init_axis_profile.argtypes = [vp, C.c_int, C.c_int, dp]
    L.dlbc_cuda_site_function.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, dp, dp, dp]
    # device pointers as integers (tensor.data_ptr()); stream handle; message buffer
    L.dlbc_cuda_mass_momentum_no_alloc.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                                   C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    for name in SYMBOLS:
        fn = getattr(L, name)
        if name not in ("dlbc_cuda_last_error", "dlbc_cuda_timestep"):
            fn.restype = C.c_int
    if L.dlbc_cuda_abi_version() != ABI_VERSION:
        raise DlbcCudaError("libdlbc_cuda.so ABI version mismatch: rebuild with `python -m dlbc_b200.build`")
    _lib = L
    return L


def device_count() -> int:
    n = C.c_int(0)
    load().dlbc_cuda_device_count(C.byref(n))
    return n.value


def nccl_unique_id() -> bytes:
    buf = (C.c_uint8 * NCCL_ID_BYTES)()
    rc = load().dlbc_cuda_nccl_unique_id(buf)
    if rc:
        raise DlbcCudaError(load().dlbc_cuda_last_error(None).decode())
    return bytes(buf)


def tables():
    """The device's copy of the connectivity tables (connectivity.d:401-418, :198-212, :253-257)."""
    import numpy as np
    v = np.empty((19, 3), dtype=np.int64)
    b = np.empty(19, dtype=np.int64)
    w = np.empty(19)
    css = C.c_double()
    load().dlbc_cuda_tables(v.ctypes.data_as(C.POINTER(C.c_int64)), b.ctypes.data_as(C.POINTER(C.c_int64)),
                            w.ctypes.data_as(C.POINTER(C.c_double)), C.byref(css))
    return v, b, w, css.value


def site_function(fn, pops_or_rho, form=0, extra=None, device=0):
    """lb_site.cuh's per-site functions on the device: fn in "density", "velocity", "psi", "eqdist" (dlbc_cuda_site_function)."""
    import numpy as np
    code = {"density": 0, "velocity": 1, "psi": 2, "eqdist": 3}[fn]
    a = np.ascontiguousarray(np.atleast_1d(np.asarray(pops_or_rho, dtype=np.float64)))
    n = a.size if code == 2 else a.size // 19
    if isinstance(form, str):
        form = (PSI if code == 2 else EQ)[form]
    out = np.empty({0: (n,), 1: (n, 3), 2: (n,), 3: (n, 19)}[code])
    dp = C.POINTER(C.c_double)
    e = None if extra is None else np.ascontiguousarray(np.asarray(extra, dtype=np.float64))
    rc = load().dlbc_cuda_site_function(device, code, int(form), n, a.ctypes.data_as(dp),
                                        None if e is None else e.ctypes.data_as(dp), out.ctypes.data_as(dp))
    if rc:
        raise DlbcCudaError(load().dlbc_cuda_last_error(None).decode())
    return out
