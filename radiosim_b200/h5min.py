"""Minimal HDF5 reader/writer for the ``.h5`` layout radiosim writes.

No HDF5 library exists on this image (``h5py``/``libhdf5`` absent, no network), so
the subset h5py's defaults produce for radiosim's files is restated here:
superblock v0, old-style groups (symbol table = B-tree ``TREE`` + local ``HEAP`` +
``SNOD``), version-1 object headers, contiguous datasets.  Covered datatypes are
the ones the reference writes (``src/radiosim/utils/utils.py:344-383``,
``src/radiosim/simulations/ppdisks.py:84-89``):

* ``<f4``, ``<f8``, ``<i8``/``<i4``/``|u1`` arrays  (``x``, ``y``, ``galaxies``, ``RA`` ...)
* fixed-length strings ``S<n>`` (``date``: ``h5py.string_dtype('utf-8', 10)``)
* scalar variable-length UTF-8 strings (ppdisk ``metadata`` / ``params``), stored
  through a global heap collection (``GCOL``).

The writer lays a file out the way libhdf5 does for h5py's defaults: writing the reference
fixture's dataset back gives the fixture **byte for byte** (``tests/test_h5_format.py``), and
every written structure is walked by an independent checker of the format specification in
the same test module.  No HDF5 library exists in the build container or on the GPU box, so
"opens in h5py" itself could not be run.  Files h5py wrote with default settings (e.g. the
reference fixture ``tests/ppdisks_128.h5``) read back here.
"""

from __future__ import annotations

import struct

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


def _pad8(n: int) -> int:
    return (n + 7) & ~7


# --------------------------------------------------------------------------- writer


def _dtype_message(dt: np.dtype) -> bytes:
    """Datatype message body (version 1)."""
    if dt.kind == "f":
        size = dt.itemsize
        if size == 4:
            props = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
            bits = (0x20, 31, 0)
        elif size == 8:
            props = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
            bits = (0x20, 63, 0)
        else:
            raise TypeError(dt)
        return struct.pack("<BBBBI", 0x11, bits[0], bits[1], bits[2], size) + props
    if dt.kind in "iu":
        size = dt.itemsize
        b0 = 0x08 if dt.kind == "i" else 0x00
        return struct.pack("<BBBBI", 0x10, b0, 0, 0, size) + struct.pack("<HH", 0, size * 8)
    if dt.kind == "S":
        # class 3 string: null-pad (1), UTF-8 (1<<4)
        return struct.pack("<BBBBI", 0x13, 0x11, 0, 0, dt.itemsize)
    raise TypeError(f"unsupported dtype {dt}")


def _vlen_str_dtype_message() -> bytes:
    # class 9 vlen, type=string(1), pad null-term(0), cset utf8(1).  The parent type libhdf5
    # stores for a variable-length string is a 1-byte unsigned integer (H5T__set_size copies
    # H5T_NATIVE_UCHAR): class 0, little-endian, unsigned, offset 0, precision 8
    base = struct.pack("<BBBBI", 0x10, 0x00, 0, 0, 1) + struct.pack("<HH", 0, 8)
    return struct.pack("<BBBBI", 0x19, 0x01, 0x01, 0, 16) + base


def _dataspace_message(shape) -> bytes:
    rank = len(shape)
    body = struct.pack("<BBBB4x", 1, rank, 1 if rank else 0, 0)
    for s in shape:
        body += struct.pack("<Q", s)
    for s in shape:
        body += struct.pack("<Q", s)
    return body


def _message(mtype: int, body: bytes, flags: int = 0) -> bytes:
    body = body + b"\x00" * (_pad8(len(body)) - len(body))
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _object_header(messages) -> bytes:
    blob = b"".join(messages)
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(blob)) + blob


def write_h5(path, datasets: dict) -> None:
    """Write ``{name: ndarray | str}`` as datasets of the root group."""
    names = sorted(datasets)
    if len(names) > 8:
        raise ValueError("h5min writes at most 8 datasets per file")
    # ---- local heap data: "" at 0, then names.  libhdf5 sizes a fresh group's heap for four
    # 8-character names (8 + 4 * 16 + 16 = 88 bytes, H5G__obj_create's default hint); keep that
    # size when the names fit, so that small files have the layout h5py gives them.
    heap = bytearray(b"\x00" * 8)
    name_off = {}
    for n in names:
        name_off[n] = len(heap)
        enc = n.encode() + b"\x00"
        heap += enc + b"\x00" * (_pad8(len(enc)) - len(enc))
    heap_data_size = max(88, _pad8(len(heap) + 16))
    free_off = len(heap)
    heap += b"\x00" * (heap_data_size - len(heap))
    struct.pack_into("<QQ", heap, free_off, 1, heap_data_size - free_off)

    # ---- fixed layout of the metadata block, in the order libhdf5 allocates it: root header,
    # B-tree node, local heap, [first dataset's object header], symbol node, [other headers]
    off_root_hdr = 0x60
    root_hdr_len = 16 + 8 + 16
    off_tree = _pad8(off_root_hdr + root_hdr_len)
    tree_len = 24 + 8 * (2 * 16 + 1) + 8 * (2 * 16)
    off_heap = off_tree + tree_len
    off_heap_data = off_heap + 32
    snod_len = 8 + 8 * 40
    cursor = off_heap_data + heap_data_size

    # ---- datasets: object headers then raw data
    headers, raws, gcols = {}, {}, {}
    for n in names:
        v = datasets[n]
        if isinstance(v, (str, bytes)):
            raw = v.encode() if isinstance(v, str) else v
            headers[n] = ("vlen", (), None)
            raws[n] = raw
        else:
            a = np.ascontiguousarray(v)
            if a.dtype.kind == "U":
                a = np.char.encode(a, "utf-8")
            if a.dtype.byteorder == ">":
                a = a.astype(a.dtype.newbyteorder("<"))
            headers[n] = ("array", a.shape, a.dtype)
            raws[n] = memoryview(a.reshape(-1).view(np.uint8)) if a.size else b""  # no copy: written straight from the array
    hdr_off, data_off, gcol_off = {}, {}, {}
    hdr_len = {}
    off_snod = None
    for n in names:
        kind, shape, dt = headers[n]
        hdr_off[n] = cursor
        rank = len(shape)
        ds_len = 8 + _pad8(8 + 16 * rank)
        dt_body = _vlen_str_dtype_message() if kind == "vlen" else _dtype_message(dt)
        used = ds_len + 8 + _pad8(len(dt_body)) + 8 + 8 + 8 + 24
        # libhdf5 gives a dataset's header 256 bytes of message space (H5D_MINHDR_SIZE) and
        # fills what is left with one NIL message
        hdr_len[n] = 16 + (256 if used + 8 <= 256 or used == 256 else used)
        cursor = _pad8(cursor + hdr_len[n])
        if off_snod is None:  # the symbol node is created when the first link is inserted
            off_snod = cursor
            cursor += snod_len
    if off_snod is None:
        off_snod = cursor
        cursor += snod_len
    cursor = (cursor + 0x7FF) & ~0x7FF if cursor <= 0x800 else _pad8(cursor)
    head_len = cursor  # superblock, group structures and object headers; the data follows
    for n in names:
        if headers[n][0] == "vlen":
            gcol_off[n] = cursor
            obj = 16 + _pad8(len(raws[n]))
            size = max(4096, _pad8(16 + obj + 16))
            gcols[n] = size
            cursor += size
            data_off[n] = cursor
            cursor += 16
        else:
            data_off[n] = cursor if len(raws[n]) else _UNDEF
            cursor = _pad8(cursor + (raws[n].nbytes if isinstance(raws[n], memoryview) else len(raws[n])))
    eof = cursor

    out = bytearray(head_len)
    tail = []  # (file offset, bytes-like) pieces behind the head, in file order
    # superblock v0
    out[0:8] = _SIG
    struct.pack_into("<BBBBBBBxHHI", out, 8, 0, 0, 0, 0, 0, 8, 8, 4, 16, 0)
    struct.pack_into("<QQQQ", out, 24, 0, _UNDEF, eof, _UNDEF)
    # root symbol table entry: name off 0, header addr, cache type 1, scratch = tree, heap
    struct.pack_into("<QQII", out, 56, 0, off_root_hdr, 1, 0)
    struct.pack_into("<QQ", out, 80, off_tree, off_heap)
    # root object header: one symbol-table message
    root = _object_header([_message(0x11, struct.pack("<QQ", off_tree, off_heap))])
    out[off_root_hdr : off_root_hdr + len(root)] = root
    # B-tree node, one child (the SNOD)
    struct.pack_into("<4sBBHQQ", out, off_tree, b"TREE", 0, 0, 1, _UNDEF, _UNDEF)
    last_name = name_off[names[-1]] if names else 0
    struct.pack_into("<QQQ", out, off_tree + 24, 0, off_snod, last_name)
    # local heap
    struct.pack_into("<4sB3xQQQ", out, off_heap, b"HEAP", 0, heap_data_size, free_off, off_heap_data)
    out[off_heap_data : off_heap_data + heap_data_size] = heap
    # symbol node
    struct.pack_into("<4sBxH", out, off_snod, b"SNOD", 1, len(names))
    for i, n in enumerate(names):
        struct.pack_into("<QQII16x", out, off_snod + 8 + 40 * i, name_off[n], hdr_off[n], 0, 0)
    # dataset headers + data
    for n in names:
        kind, shape, dt = headers[n]
        if kind == "vlen":
            dt_body = _vlen_str_dtype_message()
            nbytes = 16
        else:
            dt_body = _dtype_message(dt)
            nbytes = raws[n].nbytes if isinstance(raws[n], memoryview) else len(raws[n])
        msgs = [
            _message(0x01, _dataspace_message(shape)),
            _message(0x03, dt_body, flags=1),
            _message(0x05, struct.pack("<BBBBI", 2, 2, 2, 1, 0), flags=1),  # fill value: size 0
            _message(0x08, struct.pack("<BBQQ", 3, 1, data_off[n], nbytes)),
        ]
        room = hdr_len[n] - 16 - sum(len(m) for m in msgs)
        if room >= 8:
            msgs.append(_message(0x00, b"\x00" * (room - 8)))  # NIL: the rest of the header block
        hdr = _object_header(msgs)
        assert len(hdr) == hdr_len[n], (len(hdr), hdr_len[n])
        out[hdr_off[n] : hdr_off[n] + len(hdr)] = hdr
        if kind == "vlen":
            g = gcol_off[n]
            size = gcols[n]
            raw = raws[n]
            col = bytearray(size + 16)  # the collection, then the 16-byte descriptor that is the dataset's data
            struct.pack_into("<4sB3xQ", col, 0, b"GCOL", 1, size)
            struct.pack_into("<HH4xQ", col, 16, 1, 0, len(raw))
            col[32 : 32 + len(raw)] = raw
            free = 32 + _pad8(len(raw))
            struct.pack_into("<HH4xQ", col, free, 0, 0, size - free)
            struct.pack_into("<IQI", col, size, len(raw), g, 1)
            assert data_off[n] == g + size
            tail.append((g, col))
        elif len(raws[n]):
            tail.append((data_off[n], raws[n]))
    with open(path, "wb") as f:
        f.write(out)
        for off, piece in tail:
            f.seek(off)
            f.write(piece)
        if f.tell() < eof:
            f.truncate(eof)


# --------------------------------------------------------------------------- reader


class _Reader:
    def __init__(self, buf: bytes):
        self.b = buf
        if buf[:8] != _SIG:
            raise ValueError("not an HDF5 file")
        if buf[8] != 0 or buf[13] != 8 or buf[14] != 8:
            raise ValueError("h5min reads superblock v0 with 8-byte offsets only")
        (self.base,) = struct.unpack_from("<Q", buf, 24)
        self.root_hdr = struct.unpack_from("<Q", buf, 56 + 8)[0]

    # -- object header v1 -> list of (type, body)
    def messages(self, addr):
        b = self.b
        ver, _, nmsg, _, hsize = struct.unpack_from("<BBHII", b, addr)
        if ver != 1:
            raise ValueError("h5min reads version-1 object headers only")
        blocks = [(addr + 16, hsize)]
        out = []
        while blocks and len(out) < nmsg:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and len(out) < nmsg:
                mtype, msize, _flags = struct.unpack_from("<HHB", b, pos)
                body = b[pos + 8 : pos + 8 + msize]
                pos += 8 + msize
                if mtype == 0x10:  # continuation
                    blocks.append(struct.unpack_from("<QQ", body, 0))
                out.append((mtype, body))
        return out

    def group_entries(self, tree, heap):
        b = self.b
        assert b[heap : heap + 4] == b"HEAP"
        heap_data = struct.unpack_from("<Q", b, heap + 24)[0]
        entries = {}

        def walk(node):
            assert b[node : node + 4] == b"TREE"
            _ntype, level, used = struct.unpack_from("<BBH", b, node + 4)
            for i in range(used):
                child = struct.unpack_from("<Q", b, node + 24 + 8 + 16 * i)[0]
                if level > 0:
                    walk(child)
                else:
                    assert b[child : child + 4] == b"SNOD"
                    nsym = struct.unpack_from("<H", b, child + 6)[0]
                    for k in range(nsym):
                        noff, haddr = struct.unpack_from("<QQ", b, child + 8 + 40 * k)
                        s = heap_data + noff
                        name = b[s : b.index(b"\x00", s)].decode()
                        entries[name] = haddr

        walk(tree)
        return entries

    def dataset(self, addr):
        b = self.b
        shape = dt = layout = None
        for mtype, body in self.messages(addr):
            if mtype == 0x01:
                ver, rank, flags = body[0], body[1], body[2]
                off = 8 if ver == 1 else 4
                shape = struct.unpack_from("<" + "Q" * rank, body, off)
            elif mtype == 0x03:
                dt = body
            elif mtype == 0x08:
                layout = body
        if shape is None or dt is None or layout is None:
            raise ValueError("not a simple dataset")
        cls = dt[0] & 0x0F
        size = struct.unpack_from("<I", dt, 4)[0]
        if layout[0] != 3:
            raise ValueError("h5min reads layout version 3 only")
        lclass = layout[1]
        if lclass == 1:
            daddr, dsize = struct.unpack_from("<QQ", layout, 2)
            raw = b"" if daddr == _UNDEF else b[daddr : daddr + dsize]
        elif lclass == 0:
            (dsize,) = struct.unpack_from("<H", layout, 2)
            raw = layout[4 : 4 + dsize]
        else:
            raise ValueError("chunked datasets are not supported by h5min")
        if cls == 1:
            npdt = np.dtype("<f%d" % size)
        elif cls == 0:
            npdt = np.dtype(("<i%d" if dt[1] & 0x08 else "<u%d") % size)
        elif cls == 3:
            npdt = np.dtype("S%d" % size)
        elif cls == 9:
            out = []
            n = int(np.prod(shape)) if shape else 1
            for i in range(n):
                length, gaddr, idx = struct.unpack_from("<IQI", raw, 16 * i)
                out.append(self._gheap_object(gaddr, idx)[:length].decode("utf-8"))
            return out[0] if not shape else np.array(out, dtype=object).reshape(shape)
        else:
            raise ValueError(f"unsupported datatype class {cls}")
        return np.frombuffer(raw, dtype=npdt).reshape(shape).copy()

    def _gheap_object(self, gaddr, idx):
        b = self.b
        assert b[gaddr : gaddr + 4] == b"GCOL"
        size = struct.unpack_from("<Q", b, gaddr + 8)[0]
        pos = gaddr + 16
        while pos < gaddr + size:
            oidx, _ref, osize = struct.unpack_from("<HH4xQ", b, pos)
            if oidx == idx:
                return b[pos + 16 : pos + 16 + osize]
            if oidx == 0:
                break
            pos += 16 + _pad8(osize)
        raise KeyError(idx)


def read_h5(path) -> dict:
    """Read every dataset of the root group into ``{name: ndarray | str}``."""
    with open(path, "rb") as f:
        r = _Reader(f.read())
    tree = heap = None
    for mtype, body in r.messages(r.root_hdr):
        if mtype == 0x11:
            tree, heap = struct.unpack_from("<QQ", body, 0)
    if tree is None:
        raise ValueError("root group has no symbol table (new-style group?)")
    return {name: r.dataset(addr) for name, addr in r.group_entries(tree, heap).items()}
