"""Minimal pure-Python reader for the HDF5 files DLBC writes (no libhdf5 / h5py needed).

DLBC dumps every field as one dataset ``/OutArray`` (reference: src/dlbc/io/hdf5.d:125-285)
and its golden test data were repacked with shuffle+gzip (tools/hdf5-repack.sh:5).  All of
those files share one simple on-disk profile: superblock v0, object header v1, data layout
v3 (contiguous or chunked with a v1 B-tree), filter pipeline v1 (shuffle, deflate),
little-endian ``<f8`` or ``<i4`` elements.  This module reads exactly that profile.

It is host-side plumbing (golden-fixture extraction, checkpoint import); it is not on the
GPU hot path.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Error(RuntimeError):
    pass


class _File:
    def __init__(self, path):
        with open(path, "rb") as fh:
            self.b = fh.read()
        if self.b[:8] != _SIG:
            raise H5Error(f"{path}: not an HDF5 file")
        if self.b[8] != 0:
            raise H5Error(f"{path}: superblock version {self.b[8]} unsupported (need 0)")
        if self.b[13] != 8 or self.b[14] != 8:
            raise H5Error(f"{path}: only 8-byte offsets/lengths supported")
        # superblock v0: 24 bytes of fixed fields, 4 addresses (base, free, eof, driver),
        # then the root group symbol table entry at byte 56.
        self.base = self.u64(24)
        self.root = self._symtab_entry(56)

    def u16(self, o):
        return struct.unpack_from("<H", self.b, o)[0]

    def u32(self, o):
        return struct.unpack_from("<I", self.b, o)[0]

    def u64(self, o):
        return struct.unpack_from("<Q", self.b, o)[0]

    def _symtab_entry(self, o):
        name_off, ohdr, cache = self.u64(o), self.u64(o + 8), self.u32(o + 16)
        btree = heap = None
        if cache == 1:
            btree, heap = self.u64(o + 24), self.u64(o + 32)
        return dict(name_off=name_off, ohdr=ohdr, btree=btree, heap=heap)

    # -- groups ---------------------------------------------------------------------
    def _heap_data(self, heap_addr):
        if self.b[heap_addr:heap_addr + 4] != b"HEAP":
            raise H5Error("bad local heap signature")
        return self.u64(heap_addr + 24)

    def _group_entries(self, btree_addr, heap_addr, out):
        if self.b[btree_addr:btree_addr + 4] != b"TREE":
            raise H5Error("bad group B-tree signature")
        level = self.b[btree_addr + 5]
        used = self.u16(btree_addr + 6)
        o = btree_addr + 24
        heap_data = self._heap_data(heap_addr)
        for i in range(used):
            child = self.u64(o + 8 + 16 * i)
            if level > 0:
                self._group_entries(child, heap_addr, out)
                continue
            if self.b[child:child + 4] != b"SNOD":
                raise H5Error("bad symbol node signature")
            n = self.u16(child + 6)
            for j in range(n):
                e = self._symtab_entry(child + 8 + 40 * j)
                s = heap_data + e["name_off"]
                name = self.b[s:self.b.index(b"\0", s)].decode()
                out[name] = e

    def links(self, entry=None):
        entry = entry or self.root
        btree, heap = entry["btree"], entry["heap"]
        if btree is None:
            # symbol table message (0x11) in the object header
            for typ, body in self.messages(entry["ohdr"]):
                if typ == 0x11:
                    btree, heap = struct.unpack_from("<QQ", body, 0)
            if btree is None:
                raise H5Error("object is not an old-style group")
        out = {}
        self._group_entries(btree, heap, out)
        return out

    # -- object headers -------------------------------------------------------------
    def messages(self, addr):
        if self.b[addr] != 1:
            raise H5Error(f"object header version {self.b[addr]} unsupported (need 1)")
        nmsg = self.u16(addr + 2)
        size = self.u32(addr + 8)
        blocks = [(addr + 16, size)]
        msgs = []
        while blocks and len(msgs) < nmsg:
            o, sz = blocks.pop(0)
            end = o + sz
            while o + 8 <= end and len(msgs) < nmsg:
                typ, msz = self.u16(o), self.u16(o + 2)
                body = self.b[o + 8:o + 8 + msz]
                if typ == 0x10:
                    blocks.append(struct.unpack_from("<QQ", body, 0))
                msgs.append((typ, body))
                o += 8 + msz
        return msgs

    # -- datasets -------------------------------------------------------------------
    def dataset(self, entry):
        shape = dtype = layout = None
        filters = []
        for typ, body in self.messages(entry["ohdr"]):
            if typ == 0x01:
                if body[0] != 1:
                    raise H5Error("dataspace version != 1")
                rank = body[1]
                shape = struct.unpack_from(f"<{rank}Q", body, 8)
            elif typ == 0x03:
                cls, size = body[0] & 0xF, struct.unpack_from("<I", body, 4)[0]
                if cls == 1 and size == 8:
                    dtype = np.dtype("<f8")
                elif cls == 0 and size == 4:
                    dtype = np.dtype("<i4")
                elif cls == 0 and size == 8:
                    dtype = np.dtype("<i8")
                else:
                    raise H5Error(f"datatype class {cls} size {size} unsupported")
            elif typ == 0x08:
                if body[0] != 3:
                    raise H5Error("layout version != 3")
                if body[1] == 1:
                    layout = ("contiguous",) + struct.unpack_from("<QQ", body, 2)
                elif body[1] == 2:
                    nd = body[2]
                    bt = struct.unpack_from("<Q", body, 3)[0]
                    cdims = struct.unpack_from(f"<{nd}I", body, 11)
                    layout = ("chunked", bt, cdims)
                elif body[1] == 0:
                    size = struct.unpack_from("<H", body, 2)[0]
                    layout = ("compact", bytes(body[4:4 + size]))
                else:
                    raise H5Error("unknown layout class")
            elif typ == 0x0B:
                if body[0] != 1:
                    raise H5Error("filter pipeline version != 1")
                o = 8
                for _ in range(body[1]):
                    fid, nlen, _flags, ncv = struct.unpack_from("<HHHH", body, o)
                    o += 8 + ((nlen + 7) // 8) * 8 + 4 * ncv + (4 if ncv % 2 else 0)
                    filters.append(fid)
        if shape is None or dtype is None or layout is None:
            raise H5Error("incomplete dataset header")
        n = int(np.prod(shape)) if shape else 1
        if layout[0] == "contiguous":
            _, addr, size = layout
            if addr == _UNDEF:
                return np.zeros(shape, dtype)
            return np.frombuffer(self.b, dtype, n, addr).reshape(shape).copy()
        if layout[0] == "compact":
            return np.frombuffer(layout[1], dtype, n).reshape(shape).copy()
        _, bt, cdims = layout
        out = np.zeros(shape, dtype)
        self._read_chunks(bt, cdims[:-1], dtype, filters, out)
        return out

    def _read_chunks(self, addr, cshape, dtype, filters, out):
        if self.b[addr:addr + 4] != b"TREE" or self.b[addr + 4] != 1:
            raise H5Error("bad chunk B-tree node")
        level, used = self.b[addr + 5], self.u16(addr + 6)
        nd = len(cshape)
        keysz = 8 + 8 * (nd + 1)
        o = addr + 24
        for _ in range(used):
            csize, fmask = self.u32(o), self.u32(o + 4)
            offs = struct.unpack_from(f"<{nd}Q", self.b, o + 8)
            child = self.u64(o + keysz)
            o += keysz + 8
            if level > 0:
                self._read_chunks(child, cshape, dtype, filters, out)
                continue
            raw = self.b[child:child + csize]
            for k, fid in reversed(list(enumerate(filters))):
                if fmask & (1 << k):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    es = dtype.itemsize
                    raw = np.frombuffer(raw, np.uint8).reshape(es, -1).T.tobytes()
                else:
                    raise H5Error(f"filter id {fid} unsupported")
            chunk = np.frombuffer(raw, dtype, int(np.prod(cshape))).reshape(cshape)
            sl_out, sl_in = [], []
            for a, c, s in zip(offs, cshape, out.shape):
                e = min(a + c, s)
                sl_out.append(slice(a, e))
                sl_in.append(slice(0, e - a))
            out[tuple(sl_out)] = chunk[tuple(sl_in)]


def read_dataset(path, name="OutArray"):
    """Return dataset ``/<name>`` of an HDF5 file as a numpy array (C order)."""
    f = _File(path)
    links = f.links()
    if name not in links:
        raise H5Error(f"{path}: no dataset /{name} (have {sorted(links)})")
    return f.dataset(links[name])


def list_root(path):
    return sorted(_File(path).links())


# ---- writer ---------------------------------------------------------------------------------------
# The profile the reference writes with writeChunked = false (src/dlbc/io/hdf5.d:125-285): one contiguous
# dataset /OutArray in an old-style root group.  Superblock v0, object headers v1, dataspace v1, data layout
# v3 -- the same on-disk structures the reference's own files use (read back by this module's reader and
# laid out like the libhdf5-written golden files; there is no libhdf5 in this image to cross-check with).
# Not written: the "input"/"metadata"/"globals" attribute groups of hdf5.d:453-548 (h5diff on /OutArray,
# which is what the reference's test harness runs, does not look at them).

def _pad8(b):
    return b + b"\0" * (-len(b) % 8)


def _msg(typ, body):
    body = _pad8(body)
    return struct.pack("<HHB3x", typ, len(body), 0) + body


def _ohdr(msgs):
    data = b"".join(msgs)
    return struct.pack("<BxHII4x", 1, len(msgs), 1, len(data)) + data


def _datatype(dtype):
    dtype = np.dtype(dtype)
    if dtype == np.dtype("<f8"):  # IEEE binary64, little-endian, implied mantissa msb, sign bit 63
        return struct.pack("<B3BI", 0x11, 0x20, 0x3F, 0x00, 8) + struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
    if dtype == np.dtype("<i4"):  # two's complement, little-endian (H5T_NATIVE_INT, hdf5.d:90-92)
        return struct.pack("<B3BI", 0x10, 0x08, 0x00, 0x00, 4) + struct.pack("<HH", 0, 32)
    raise H5Error(f"dtype {dtype} unsupported (need <f8 or <i4)")


def write_dataset(path, array, name="OutArray"):
    """Write `array` (float64 or int32, C order) as the contiguous dataset /<name> of a new HDF5 file."""
    a = np.ascontiguousarray(array)
    if a.dtype == np.float64:
        a = a.astype("<f8", copy=False)
    elif a.dtype == np.int32:
        a = a.astype("<i4", copy=False)
    else:
        raise H5Error(f"dtype {a.dtype} unsupported (need float64 or int32)")
    K_LEAF, K_INT = 4, 16
    nm = name.encode() + b"\0"
    heap_data = _pad8(b"\0") + _pad8(nm)                       # offset 0: "", offset 8: name
    name_off = 8
    o_root = 96
    root = _ohdr([_msg(0x11, struct.pack("<QQ", 0, 0))])        # patched below
    o_btree = o_root + len(root)
    btree_size = 24 + (2 * K_INT + 1) * 8 + 2 * K_INT * 8
    o_heap = o_btree + btree_size
    o_heap_data = o_heap + 32
    o_snod = o_heap_data + len(heap_data)
    snod_size = 8 + 2 * K_LEAF * 40
    o_dset = o_snod + snod_size
    rank = a.ndim
    dataspace = struct.pack("<BBBx4x", 1, rank, 0) + struct.pack(f"<{rank}Q", *a.shape)
    fill = bytes([2, 2, 2, 1, 0, 0, 0, 0])                      # v2: late allocation, fill if set, default value
    layout_len = len(_msg(0x08, struct.pack("<BBQQ", 3, 1, 0, 0)))
    dset_len = 16 + len(_msg(0x01, dataspace)) + len(_msg(0x03, _datatype(a.dtype))) + len(_msg(0x05, fill)) + layout_len
    o_data = (o_dset + dset_len + 7) // 8 * 8
    eof = o_data + a.nbytes
    dset = _ohdr([_msg(0x01, dataspace), _msg(0x03, _datatype(a.dtype)), _msg(0x05, fill),
                  _msg(0x08, struct.pack("<BBQQ", 3, 1, o_data, a.nbytes))])
    assert len(dset) == dset_len
    root = _ohdr([_msg(0x11, struct.pack("<QQ", o_btree, o_heap))])
    sb = _SIG + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", K_LEAF, K_INT, 0)
    sb += struct.pack("<QQQQ", 0, _UNDEF, eof, _UNDEF)
    sb += struct.pack("<QQI4xQQ", 0, o_root, 1, o_btree, o_heap)  # root group symbol table entry
    assert len(sb) == 96
    btree = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, _UNDEF, _UNDEF) + struct.pack("<QQQ", 0, o_snod, name_off)
    btree += b"\0" * (btree_size - len(btree))
    heap = b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), 1, o_heap_data)  # free-list head 1 = none
    snod = b"SNOD" + struct.pack("<BxH", 1, 1) + struct.pack("<QQI4x16x", name_off, o_dset, 0)
    snod += b"\0" * (snod_size - len(snod))
    with open(path, "wb") as fh:
        fh.write(sb + root + btree + heap + heap_data + snod + dset)
        fh.write(b"\0" * (o_data - (o_dset + dset_len)))
        fh.write(a.tobytes())
    return path


def field_file_name(field, simulation_name, timestep, time_id="00000000T000000"):
    """makeFilenameOutput (io/io.d:157-172): <field>-<simulationName>-<id>-t%08d.h5; the reference's id is
    the start time of the run (yyyymmddThhmmss)."""
    return f"{field}-{simulation_name}-{time_id}-t{int(timestep):08d}.h5"
