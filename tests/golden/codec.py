"""Fixture codec: float64/int32 arrays stored byte-shuffled inside a compressed .npz.

Byte-shuffling (all first bytes, then all second bytes, ...) is what HDF5's shuffle filter does
and makes smooth fp64 fields compress about as well as the reference's own golden files.
Values round-trip bit-exactly.
"""
import numpy as np


def save(path, **arrays):
    packed = {}
    for name, a in arrays.items():
        a = np.ascontiguousarray(a)
        packed[name + "__shape"] = np.array(a.shape, dtype=np.int64)
        packed[name + "__dtype"] = np.array(a.dtype.str)
        packed[name] = a.view(np.uint8).reshape(-1, a.dtype.itemsize).T.copy()
    np.savez_compressed(path, **packed)


def load(path):
    out = {}
    with np.load(path) as z:
        for name in z.files:
            if name.endswith("__shape") or name.endswith("__dtype"):
                continue
            dt = np.dtype(str(z[name + "__dtype"]))
            shape = tuple(int(x) for x in z[name + "__shape"])
            out[name] = np.ascontiguousarray(z[name].T).view(dt).reshape(shape)
    return out


def digest(a):
    """SHA-256 of an array's shape, dtype and bytes, with every NaN replaced by the canonical quiet NaN (a field that
    was never initialised holds D's double.init; NaN payloads and signs are not part of what the goldens pin)."""
    import hashlib
    a = np.ascontiguousarray(a)
    if a.dtype.kind == "f":
        a = np.where(np.isnan(a), np.float64("nan"), a.astype(np.float64))
    h = hashlib.sha256()
    h.update(repr((a.shape, a.dtype.str)).encode())
    h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()
