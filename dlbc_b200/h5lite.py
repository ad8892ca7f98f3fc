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


def _parse_attribute(f, body):
    """Attribute message v1 (the only version the reference's files carry): name, datatype, dataspace, data."""
    if body[0] != 1:
        raise H5Error(f"attribute message version {body[0]} unsupported (need 1)")
    nsz, tsz, ssz = struct.unpack_from("<HHH", body, 2)
    o = 8
    name = bytes(body[o:o + nsz]).split(b"\0")[0].decode()
    o += (nsz + 7) // 8 * 8
    dt = body[o:o + tsz]
    o += (tsz + 7) // 8 * 8
    ds = body[o:o + ssz]
    o += (ssz + 7) // 8 * 8
    rank = ds[1]
    dims = struct.unpack_from(f"<{rank}Q", ds, 8) if rank else ()
    n = int(np.prod(dims)) if rank else 1
    cls, size = dt[0] & 0xF, struct.unpack_from("<I", dt, 4)[0]
    if cls == 9 and (dt[1] & 0xF) == 1:  # variable-length string: (length u32, global heap collection u64, index u32)
        vals = []
        for i in range(n):
            ln, addr, idx = struct.unpack_from("<IQI", body, o + 16 * i)
            vals.append(f._gheap_object(addr, idx)[:ln].split(b"\0")[0].decode() if addr not in (0, _UNDEF) else "")
        return name, vals
    if cls == 0 and size in (1, 2, 4, 8):
        signed = bool(dt[1] & 0x08)
        code = {1: "b", 2: "h", 4: "i", 8: "q"}[size]
        vals = list(struct.unpack_from("<" + str(n) + (code if signed else code.upper()), body, o))
        return name, vals[0] if n == 1 else vals
    if cls == 1 and size == 8:
        vals = list(struct.unpack_from(f"<{n}d", body, o))
        return name, vals[0] if n == 1 else vals
    raise H5Error(f"attribute {name}: datatype class {cls} size {size} unsupported")


def _gheap_object(self, addr, index):
    """Object `index` of the global heap collection at `addr` (variable-length data lives there)."""
    if self.b[addr:addr + 4] != b"GCOL":
        raise H5Error("bad global heap signature")
    end = addr + self.u64(addr + 8)
    o = addr + 16
    while o + 16 <= end:
        idx, size = self.u16(o), self.u64(o + 8)
        if idx == index:
            return self.b[o + 16:o + 16 + size]
        if idx == 0:
            break
        o += 16 + (size + 7) // 8 * 8
    raise H5Error(f"global heap object {index} not found")


_File._gheap_object = _gheap_object


def read_attributes(path, group=None):
    """Attributes of the root group (group=None) or of the group /<group> as {name: value}: variable-length string
    attributes become lists of str (the reference's `input` attribute: the lines of the parameter file,
    io/hdf5.d:529-548; `metadata/revision*`, :290-299), integers / doubles scalars (`globals/dlbc.lb.lb.timestep`,
    io/checkpoint.d:248-313)."""
    f = _File(path)
    entry = f.root
    if group is not None:
        links = f.links()
        if group not in links:
            raise H5Error(f"{path}: no group /{group} (have {sorted(links)})")
        entry = links[group]
    out = {}
    for typ, body in f.messages(entry["ohdr"]):
        if typ == 0x0C:
            k, v = _parse_attribute(f, body)
            out[k] = v
    return out


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
# Also written (write_file): the `input` attribute, the `metadata` group and, for checkpoints, the `globals`
# group of hdf5.d:264-299,453-548 and checkpoint.d:248-313, in the encodings found in the reference's own files.

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


_VLEN_STR = struct.pack("<B3BI", 0x19, 0x01, 0x00, 0x00, 16) + struct.pack("<B3BIHH", 0x10, 0, 0, 0, 1, 0, 8)  # as libhdf5 encodes it
_INT32 = struct.pack("<B3BI", 0x10, 0x08, 0x00, 0x00, 4) + struct.pack("<HH", 0, 32)
_SPACE_1D = lambda n: struct.pack("<BBBx4x", 1, 1, 1) + struct.pack("<QQ", n, n)  # noqa: E731  (v1, max dims present)


class _GlobalHeap:
    """One global heap collection: the variable-length strings of every attribute of the file."""

    def __init__(self):
        self.objs = []

    def add(self, data):
        self.objs.append(data)
        return len(self.objs)  # indices start at 1

    def size(self):
        n = 16 + sum(16 + (len(d) + 7) // 8 * 8 for d in self.objs) + 16  # header, objects, the free-space object 0
        return max(4096, (n + 7) // 8 * 8)

    def bytes(self):
        body = b""
        for i, d in enumerate(self.objs):
            body += struct.pack("<HH4xQ", i + 1, 0, len(d)) + _pad8(d)
        total = self.size()
        free = total - 16 - len(body)
        body += struct.pack("<HH4xQ", 0, 0, free) + b"\0" * (free - 16)
        return b"GCOL" + struct.pack("<B3xQ", 1, total) + body


def _attr_msg(name, datatype, nelem, data):
    nm = name.encode() + b"\0"
    space = _SPACE_1D(nelem)
    body = struct.pack("<BxHHH", 1, len(nm), len(datatype), len(space)) + _pad8(nm) + _pad8(datatype) + _pad8(space) + data
    if len(body) > 65528:
        raise H5Error(f"attribute {name} does not fit an object header message ({len(body)} bytes)")
    return _msg(0x0C, body)


def _string_attr(name, strings, gheap, gheap_addr_slot):
    """vlen-string attribute; the heap address is not known yet: `gheap_addr_slot` collects (message, offsets) to patch."""
    data = b""
    for st in strings:
        raw = st.encode() + b"\0"
        data += struct.pack("<IQI", len(raw), 0, gheap.add(raw))
    m = bytearray(_attr_msg(name, _VLEN_STR, len(strings), data))
    first = len(m) - len(_pad8(data))
    gheap_addr_slot.append((m, [first + 16 * i + 4 for i in range(len(strings))]))
    return m


_CHUNK_K = 32  # indexed-storage internal node K of a version-0 superblock (fixed: the field only exists from v1 on)


def _chunk_index(a, chunks, o_tree):
    """Chunked storage of `a` (io/hdf5.d:237-241: `writeChunked = true` makes every rank's block one chunk): returns
    (tree bytes, [chunk bytes ...]) for a version-1 B-tree of node type 1 placed at file offset o_tree with the raw
    chunks stored back to back right after it.  Chunks are whole (edge chunks are padded with zeros, as libhdf5 stores
    them), unfiltered, keyed by their element offsets + a trailing 0 for the element-size dimension; the key after the
    last one of a node is `last chunk offset + chunk dims` as in the files libhdf5 writes.  One leaf holds up to
    2K = 64 chunks; more chunks get one internal node above up to 64 leaves."""
    nd, es = a.ndim, a.dtype.itemsize
    if len(chunks) != nd or any(int(c) < 1 for c in chunks):
        raise H5Error(f"chunk dims {chunks} do not fit an array of rank {nd}")
    chunks = tuple(int(c) for c in chunks)
    grid = [(s + c - 1) // c for s, c in zip(a.shape, chunks)]
    offs_list = [tuple(int(i) * c for i, c in zip(idx, chunks)) for idx in np.ndindex(*grid)]
    cap = 2 * _CHUNK_K
    if len(offs_list) > cap * cap:
        raise H5Error(f"{len(offs_list)} chunks: more than a two-level index holds ({cap * cap})")
    csize = int(np.prod(chunks)) * es
    keysz = 8 + 8 * (nd + 1)
    node_size = 24 + (cap + 1) * keysz + cap * 8
    leaves = [offs_list[i:i + cap] for i in range(0, len(offs_list), cap)]
    n_nodes = len(leaves) + (1 if len(leaves) > 1 else 0)
    o_chunks = o_tree + n_nodes * node_size

    def key(size, offs):
        return struct.pack("<II", size, 0) + struct.pack(f"<{nd + 1}Q", *offs, 0)

    def end_key(offs):
        return struct.pack("<II", 0, 0) + struct.pack(f"<{nd + 1}Q", *[o + c for o, c in zip(offs, chunks)], es)

    def node(level, entries, left, right):  # entries: [(offsets, child address)]
        b = b"TREE" + struct.pack("<BBHQQ", 1, level, len(entries), left, right)
        for offs, child in entries:
            b += key(csize, offs) + struct.pack("<Q", child)
        b += end_key(entries[-1][0])
        return b + b"\0" * (node_size - len(b))

    first_leaf = o_tree + (node_size if len(leaves) > 1 else 0)
    tree, k = b"", 0
    for i, lf in enumerate(leaves):
        left = first_leaf + (i - 1) * node_size if i > 0 else _UNDEF
        right = first_leaf + (i + 1) * node_size if i + 1 < len(leaves) else _UNDEF
        tree += node(0, [(o, o_chunks + (k + j) * csize) for j, o in enumerate(lf)], left, right)
        k += len(lf)
    if len(leaves) > 1:
        tree = node(1, [(lf[0], first_leaf + i * node_size) for i, lf in enumerate(leaves)], _UNDEF, _UNDEF) + tree
    data = []
    for offs in offs_list:
        blk = a[tuple(slice(o, o + c) for o, c in zip(offs, chunks))]
        if blk.shape != chunks:
            full = np.zeros(chunks, a.dtype)
            full[tuple(slice(0, n) for n in blk.shape)] = blk
            blk = full
        data.append(np.ascontiguousarray(blk).tobytes())
    return tree, data


def write_file(path, array, name="OutArray", input_lines=None, metadata=None, globals_=None, chunks=None):
    """Write `array` (float64 or int32, C order) as the dataset /<name> of a new HDF5 file -- contiguous, or chunked
    when `chunks` (chunk dims, one per array dimension) is given: the reference's `hdf5.writeChunked = true` form,
    where the chunk is one rank's block (io/hdf5.d:237-241; replay.write_outputs passes the z-slab) --, with the
    attributes the reference adds to every field file (io/hdf5.d:264-285): `input` on the root group -- the lines of
    the parameter file, from which the reference can re-read its parameters (parameters.d:172-174,339-351) --, the
    group `metadata` with string / int attributes (revision info, hdf5.d:290-299) and, for checkpoint files, the
    group `globals` with int attributes (`dlbc.lb.lb.timestep`, io/checkpoint.d:248-313).
    Same structures, byte for byte where it matters, as the libhdf5-written files under the reference's
    reference-data (superblock v0, object header v1, attribute message v1, variable-length strings in a global
    heap collection, old-style groups with a B-tree and a local heap)."""
    a = np.ascontiguousarray(array)
    if a.dtype == np.float64:
        a = a.astype("<f8", copy=False)
    elif a.dtype == np.int32:
        a = a.astype("<i4", copy=False)
    else:
        raise H5Error(f"dtype {a.dtype} unsupported (need float64 or int32)")
    K_LEAF, K_INT = 4, 16
    gheap, patches = _GlobalHeap(), []
    groups = {}
    if metadata is not None:
        groups["metadata"] = metadata
    if globals_ is not None:
        groups["globals"] = globals_
    names = sorted([name] + list(groups))  # symbol node entries are sorted by name
    heap_data, name_off = _pad8(b"\0"), {}
    for nm in names:
        name_off[nm] = len(heap_data)
        heap_data += _pad8(nm.encode() + b"\0")
    btree_size = 24 + (2 * K_INT + 1) * 8 + 2 * K_INT * 8
    snod_size = 8 + 2 * K_LEAF * 40

    def attr_msgs(attrs):
        out = []
        for k, v in attrs.items():
            if isinstance(v, str):
                out.append(_string_attr(k, [v], gheap, patches))
            elif isinstance(v, (list, tuple)) and all(isinstance(x, str) for x in v):
                out.append(_string_attr(k, list(v), gheap, patches))
            elif isinstance(v, (bool, int, np.integer)):
                out.append(bytearray(_attr_msg(k, _INT32, 1, struct.pack("<i", int(v)))))
            else:
                raise H5Error(f"attribute {k}: unsupported value {v!r}")
        return out

    root_attrs = attr_msgs({"input": list(input_lines)} if input_lines is not None else {})
    group_attrs = {g: attr_msgs(at) for g, at in groups.items()}
    # ---- layout --------------------------------------------------------------------------------------------
    o = 96
    o_root = o
    root_len = 16 + 24 + sum(len(m) for m in root_attrs)
    o += root_len
    o_btree = o; o += btree_size
    o_heap = o; o += 32
    o_heap_data = o; o += len(heap_data)
    o_snod = o; o += snod_size
    gpos = {}
    for g in groups:
        glen = 16 + 24 + sum(len(m) for m in group_attrs[g])
        gpos[g] = dict(ohdr=o, btree=o + glen, heap=o + glen + btree_size, heap_data=o + glen + btree_size + 32)
        o += glen + btree_size + 32 + 8
    o_dset = o
    rank = a.ndim
    dataspace = struct.pack("<BBBx4x", 1, rank, 0) + struct.pack(f"<{rank}Q", *a.shape)
    fill = bytes([2, 2, 2, 1, 0, 0, 0, 0])                      # v2: late allocation, fill if set, default value
    if chunks is None:
        layout0 = struct.pack("<BBQQ", 3, 1, 0, 0)
    else:
        layout0 = struct.pack("<BBBQ", 3, 2, rank + 1, 0) + struct.pack(f"<{rank + 1}I", *([1] * rank), a.dtype.itemsize)
    dset_len = 16 + len(_msg(0x01, dataspace)) + len(_msg(0x03, _datatype(a.dtype))) + len(_msg(0x05, fill)) + \
        len(_msg(0x08, layout0))
    o += dset_len
    o_gheap = (o + 7) // 8 * 8
    o = o_gheap + (gheap.size() if gheap.objs else 0)
    o_data = (o + 7) // 8 * 8
    if chunks is None:
        payload = [a.tobytes()]
        layout = struct.pack("<BBQQ", 3, 1, o_data, a.nbytes)
    else:
        tree, payload = _chunk_index(a, chunks, o_data)
        payload = [tree] + payload
        layout = struct.pack("<BBBQ", 3, 2, rank + 1, o_data) + struct.pack(f"<{rank + 1}I", *[int(c) for c in chunks],
                                                                            a.dtype.itemsize)
    eof = o_data + sum(len(b) for b in payload)
    for m, offs in patches:
        for k in offs:
            struct.pack_into("<Q", m, k, o_gheap)
    # ---- pieces --------------------------------------------------------------------------------------------
    sb = _SIG + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", K_LEAF, K_INT, 0)
    sb += struct.pack("<QQQQ", 0, _UNDEF, eof, _UNDEF)
    sb += struct.pack("<QQI4xQQ", 0, o_root, 1, o_btree, o_heap)  # root group symbol table entry
    assert len(sb) == 96
    root = _ohdr([_msg(0x11, struct.pack("<QQ", o_btree, o_heap))] + [bytes(m) for m in root_attrs])
    assert len(root) == root_len
    btree = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, _UNDEF, _UNDEF) + struct.pack("<QQQ", 0, o_snod, name_off[names[-1]])
    btree += b"\0" * (btree_size - len(btree))
    heap = b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), 1, o_heap_data)  # free-list head 1 = none
    snod = b"SNOD" + struct.pack("<BxH", 1, len(names))
    for nm in names:
        if nm in groups:
            snod += struct.pack("<QQI4xQQ", name_off[nm], gpos[nm]["ohdr"], 1, gpos[nm]["btree"], gpos[nm]["heap"])
        else:
            snod += struct.pack("<QQI4x16x", name_off[nm], o_dset, 0)
    snod += b"\0" * (snod_size - len(snod))
    blob = sb + root + btree + heap + heap_data + snod
    for g in groups:
        gp = gpos[g]
        gh = _ohdr([_msg(0x11, struct.pack("<QQ", gp["btree"], gp["heap"]))] + [bytes(m) for m in group_attrs[g]])
        gb = b"TREE" + struct.pack("<BBHQQ", 0, 0, 0, _UNDEF, _UNDEF) + struct.pack("<Q", 0)  # an empty leaf: no links
        gb += b"\0" * (btree_size - len(gb))
        ghp = b"HEAP" + struct.pack("<B3xQQQ", 0, 8, 1, gp["heap_data"]) + b"\0" * 8
        assert len(blob) == gp["ohdr"]
        blob += gh + gb + ghp
    assert len(blob) == o_dset
    dset = _ohdr([_msg(0x01, dataspace), _msg(0x03, _datatype(a.dtype)), _msg(0x05, fill), _msg(0x08, layout)])
    assert len(dset) == dset_len
    blob += dset
    blob += b"\0" * (o_gheap - len(blob))
    if gheap.objs:
        blob += gheap.bytes()
    blob += b"\0" * (o_data - len(blob))
    with open(path, "wb") as fh:
        fh.write(blob)
        for b in payload:
            fh.write(b)
    return path


def write_dataset(path, array, name="OutArray", chunks=None):
    """A bare dataset /<name> (no attributes), contiguous or chunked."""
    return write_file(path, array, name, chunks=chunks)


def field_file_name(field, simulation_name, timestep, time_id="00000000T000000"):
    """makeFilenameOutput (io/io.d:157-172): <field>-<simulationName>-<id>-t%08d.h5; the reference's id is
    the start time of the run (yyyymmddThhmmss)."""
    return f"{field}-{simulation_name}-{time_id}-t{int(timestep):08d}.h5"
