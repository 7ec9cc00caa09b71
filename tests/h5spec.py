"""Independent structural checker for the HDF5 subset radiosim's ``.h5`` files use.

Test infrastructure.  Written against the *HDF5 File Format Specification* (superblock
version 0, version-1 B-trees of group nodes, local heaps, symbol-table nodes, version-1 object
headers, global heap collections) and deliberately shares no code with
``radiosim_b200/h5min.py``: it re-parses a file field by field, asserts every reserved field,
alignment, size relation and ordering rule libhdf5 relies on, and returns a description of each
dataset (decoded messages, raw-data extent) that the tests compare between files.

No HDF5 library exists in the build container or on the GPU box, so this walker plus the
byte-for-byte comparison with the reference's h5py-written fixture is the external check of the
writer.
"""

import struct

UNDEF = 0xFFFFFFFFFFFFFFFF


class FormatError(AssertionError):
    pass


def need(cond, msg):
    if not cond:
        raise FormatError(msg)


def u(buf, fmt, off):
    return struct.unpack_from("<" + fmt, buf, off)


def cstring(buf, off):
    end = buf.index(b"\x00", off)
    return buf[off:end].decode("utf-8")


def parse_datatype(body, where):
    need(len(body) >= 8, f"{where}: datatype message shorter than 8 bytes")
    cv, b0, b1, b2, size = u(body, "BBBBI", 0)
    version, cls = cv >> 4, cv & 0x0F
    need(version == 1, f"{where}: datatype version {version} (h5py default writes 1 for these classes)")
    d = {"class": cls, "size": size}
    if cls == 0:  # fixed point
        need(len(body) >= 12, f"{where}: fixed-point properties missing")
        need(b0 & 0x01 == 0, f"{where}: big-endian integer")
        need(b0 & 0xF0 == 0 and b1 == 0 and b2 == 0, f"{where}: reserved integer class bits set")
        off, prec = u(body, "HH", 8)
        need(off == 0 and prec == 8 * size, f"{where}: integer offset/precision {off}/{prec} for size {size}")
        d.update(signed=bool(b0 & 0x08), used=12)
    elif cls == 1:  # floating point
        need(len(body) >= 20, f"{where}: float properties missing")
        need(b0 & 0x41 == 0, f"{where}: non-little-endian float")
        need((b0 >> 4) & 3 == 2, f"{where}: mantissa normalisation is not 'implied msb'")
        need(b0 & 0x0E == 0 and b2 == 0, f"{where}: float padding/reserved bits set")
        off, prec, eloc, esize, mloc, msize, bias = u(body, "HHBBBBI", 8)
        expect = {4: (32, 23, 8, 0, 23, 127, 31), 8: (64, 52, 11, 0, 52, 1023, 63)}.get(size)
        need(expect is not None, f"{where}: float size {size}")
        need((prec, eloc, esize, mloc, msize, bias, b1) == expect and off == 0,
             f"{where}: not IEEE-754 binary{8 * size}: {(off, prec, eloc, esize, mloc, msize, bias, b1)}")
        d.update(used=20)
    elif cls == 3:  # string
        pad, cset = b0 & 0x0F, b0 >> 4
        need(pad in (0, 1, 2) and cset in (0, 1) and b1 == 0 and b2 == 0, f"{where}: string class bits")
        d.update(strpad=pad, cset=cset, used=8)
    elif cls == 9:  # variable length
        vtype, pad, cset = b0 & 0x0F, b0 >> 4, b1 & 0x0F
        need(vtype in (0, 1) and pad in (0, 1, 2) and cset in (0, 1) and b1 >> 4 == 0 and b2 == 0,
             f"{where}: vlen class bits")
        need(size == 16, f"{where}: vlen element size {size} (4 + 8-byte address + 4)")
        base = parse_datatype(body[8:], where + " (vlen base)")
        if vtype == 1:
            need(base["class"] == 0 and base["size"] == 1 and not base["signed"],
                 f"{where}: a vlen STRING's base type is a 1-byte unsigned integer in libhdf5 files")
        d.update(vlen_type=vtype, strpad=pad, cset=cset, base=base, used=8 + base["used"])
    else:
        raise FormatError(f"{where}: datatype class {cls} is outside the subset")
    return d


def parse_object_header(buf, addr, where):
    need(addr % 8 == 0, f"{where}: object header at unaligned address {addr:#x}")
    ver, rsv, nmsg, refs, hsize = u(buf, "BBHII", addr)
    need(ver == 1 and rsv == 0, f"{where}: object header version {ver}, reserved {rsv}")
    need(refs == 1, f"{where}: reference count {refs}")
    need(hsize % 8 == 0, f"{where}: header size {hsize} not a multiple of 8")
    need(u(buf, "I", addr + 12)[0] == 0, f"{where}: padding after the header prefix not zero")
    pos, end = addr + 16, addr + 16 + hsize
    need(end <= len(buf), f"{where}: header runs past the end of the file")
    msgs = []
    while len(msgs) < nmsg:
        need(pos + 8 <= end, f"{where}: {nmsg} messages announced, block holds {len(msgs)}")
        mtype, msize, flags, r0, r1, r2 = u(buf, "HHBBBB", pos)
        need(r0 == r1 == r2 == 0, f"{where}: message {mtype:#x} reserved bytes set")
        need(msize % 8 == 0, f"{where}: message {mtype:#x} size {msize} not a multiple of 8")
        need(pos + 8 + msize <= end, f"{where}: message {mtype:#x} overruns the header block")
        need(mtype != 0x10, f"{where}: continuation blocks are outside the subset")
        msgs.append((mtype, flags, buf[pos + 8 : pos + 8 + msize]))
        pos += 8 + msize
    need(pos == end, f"{where}: {end - pos} bytes of the header block are not covered by messages")
    return msgs


def parse_dataset(buf, addr, name):
    where = f"dataset {name!r}"
    msgs = parse_object_header(buf, addr, where)
    types = [m[0] for m in msgs]
    for required in (0x01, 0x03, 0x08):
        need(types.count(required) == 1, f"{where}: message {required:#x} occurs {types.count(required)} times")
    need(types.count(0x05) == 1, f"{where}: libhdf5 >= 1.8 always writes one fill-value message (0x05)")
    d = {"name": name, "message_types": types, "message_flags": [m[1] for m in msgs]}
    for mtype, flags, body in msgs:
        if mtype == 0x00:
            need(body == b"\x00" * len(body), f"{where}: NIL message is not zero-filled")
        elif mtype == 0x01:
            ver, rank, sflags, rsv = u(body, "BBBB", 0)
            need(ver == 1 and rsv == 0 and u(body, "I", 4)[0] == 0, f"{where}: dataspace v{ver}/reserved")
            need(sflags in (0, 1), f"{where}: dataspace flags {sflags}")
            need(len(body) >= 8 + 8 * rank * (2 if sflags & 1 else 1), f"{where}: dataspace too short")
            dims = u(body, "Q" * rank, 8)
            if sflags & 1:
                need(u(body, "Q" * rank, 8 + 8 * rank) == dims, f"{where}: max dims differ from dims")
            else:
                need(rank == 0, f"{where}: h5py writes max dims for every non-scalar dataset")
            d["shape"] = tuple(dims)
        elif mtype == 0x03:
            need(flags & 1, f"{where}: datatype message not flagged constant")
            d["dtype"] = parse_datatype(body, where)
            need(body[d["dtype"]["used"]:] == b"\x00" * (len(body) - d["dtype"]["used"]),
                 f"{where}: bytes after the datatype are not zero")
        elif mtype == 0x05:
            ver, alloc, write, defined = u(body, "BBBB", 0)
            need(ver == 2, f"{where}: fill value version {ver}")
            need(alloc in (1, 2, 3) and write in (0, 1, 2) and defined in (0, 1), f"{where}: fill value fields")
            if defined:
                need(len(body) >= 8 and u(body, "I", 4)[0] == 0, f"{where}: a default fill value has size 0")
            d["fill"] = (alloc, write, defined)
        elif mtype == 0x08:
            ver, lclass = u(body, "BB", 0)
            need(ver == 3 and lclass == 1, f"{where}: layout v{ver} class {lclass} (contiguous v3 expected)")
            daddr, dsize = u(body, "QQ", 2)
            d["data"] = (daddr, dsize)
        elif mtype == 0x12:
            pass  # modification time: h5py can be told to write it; harmless
        else:
            raise FormatError(f"{where}: message type {mtype:#x} is outside the subset")
    n = 1
    for s in d["shape"]:
        n *= s
    daddr, dsize = d["data"]
    need(dsize == n * d["dtype"]["size"], f"{where}: layout size {dsize} != {n} elements x {d['dtype']['size']}")
    if dsize:
        need(daddr != UNDEF and daddr + dsize <= len(buf), f"{where}: raw data outside the file")
    return d


def parse_global_heap_object(buf, gaddr, index, where):
    need(buf[gaddr : gaddr + 4] == b"GCOL", f"{where}: no GCOL signature at {gaddr:#x}")
    ver, r0, r1, r2, size = u(buf, "BBBBQ", gaddr + 4)
    need(ver == 1 and r0 == r1 == r2 == 0, f"{where}: global heap version/reserved")
    need(size >= 4096 and size % 8 == 0 and gaddr + size <= len(buf), f"{where}: collection size {size}")
    pos, end, found, seen_free = gaddr + 16, gaddr + size, None, False
    while pos + 16 <= end:
        idx, refs, rsv, osize = u(buf, "HHIQ", pos)
        need(rsv == 0, f"{where}: global heap object reserved field")
        if idx == 0:  # free space: its size counts its own header and must reach the end
            need(osize == end - pos, f"{where}: free-space object size {osize}, {end - pos} bytes remain")
            seen_free = True
            break
        if idx == index:
            found = buf[pos + 16 : pos + 16 + osize]
        pos += 16 + ((osize + 7) & ~7)
    need(seen_free or pos == end, f"{where}: collection not terminated")
    need(found is not None, f"{where}: object {index} not in the collection")
    return found


def walk(path_or_bytes):
    """Validate a file, return ``{"eof": int, "datasets": [description, ...]}`` in link order."""
    buf = path_or_bytes if isinstance(path_or_bytes, (bytes, bytearray)) else open(path_or_bytes, "rb").read()
    need(buf[:8] == b"\x89HDF\r\n\x1a\n", "signature")
    sb, fs, rg, r0, sh, so, sl, r1 = u(buf, "BBBBBBBB", 8)
    need((sb, fs, rg, r0, sh, r1) == (0, 0, 0, 0, 0, 0), "superblock version fields / reserved bytes")
    need(so == 8 and sl == 8, "8-byte offsets and lengths")
    leaf_k, int_k, cflags = u(buf, "HHI", 16)
    need(leaf_k >= 1 and int_k >= 1 and cflags == 0, "group K values / consistency flags")
    base, freesp, eof, driver = u(buf, "QQQQ", 24)
    need(base == 0 and freesp == UNDEF and driver == UNDEF, "base / free-space / driver addresses")
    need(eof == len(buf), f"end-of-file address {eof} != file size {len(buf)}")
    name_off, root_hdr, cache, rsv = u(buf, "QQII", 56)
    need(name_off == 0 and rsv == 0, "root symbol-table entry")
    need(cache == 1, "root entry should cache the B-tree / heap addresses (cache type 1)")
    c_tree, c_heap = u(buf, "QQ", 80)
    root_msgs = parse_object_header(buf, root_hdr, "root group")
    stab = [m for m in root_msgs if m[0] == 0x11]
    need(len(stab) == 1, "root group needs exactly one symbol-table message")
    tree, heap = u(stab[0][2], "QQ", 0)
    need((tree, heap) == (c_tree, c_heap), "cached B-tree/heap addresses differ from the symbol-table message")

    # local heap
    need(buf[heap : heap + 4] == b"HEAP", "local heap signature")
    hver, h0, h1, h2 = u(buf, "BBBB", heap + 4)
    need((hver, h0, h1, h2) == (0, 0, 0, 0), "local heap version/reserved")
    hsize, hfree, hdata = u(buf, "QQQ", heap + 8)
    need(hdata % 8 == 0 and hsize % 8 == 0 and hdata + hsize <= len(buf), "local heap data segment")
    need(buf[hdata] == 0, "heap offset 0 must hold the empty string")
    free_blocks = []
    nxt = hfree
    while nxt != UNDEF and nxt != 1:
        need(nxt % 8 == 0 and nxt + 16 <= hsize, f"local heap free block at {nxt}")
        nn, fsz = u(buf, "QQ", hdata + nxt)
        need(fsz >= 16 and nxt + fsz <= hsize, f"local heap free block size {fsz}")
        free_blocks.append((nxt, fsz))
        nxt = nn

    def in_free(off):
        return any(a <= off < a + s for a, s in free_blocks)

    # B-tree of group nodes
    entries = []

    def node(addr, level_expected=None):
        need(buf[addr : addr + 4] == b"TREE", f"B-tree signature at {addr:#x}")
        ntype, level, used = u(buf, "BBH", addr + 4)
        left, right = u(buf, "QQ", addr + 8)
        need(ntype == 0, "B-tree node type 0 (group)")
        need(level_expected is None or level == level_expected, "B-tree level")
        need(1 <= used <= 2 * int_k, f"B-tree entries used {used}")
        need(left == UNDEF and right == UNDEF, "single-node levels only in this subset")
        need(addr + 24 + 8 * (2 * int_k + 1) + 8 * 2 * int_k <= len(buf), "B-tree node truncated")
        keys = [u(buf, "Q", addr + 24 + 16 * i)[0] for i in range(used + 1)]
        kids = [u(buf, "Q", addr + 24 + 16 * i + 8)[0] for i in range(used)]
        for k in keys:
            need(k < hsize and not in_free(k), f"B-tree key {k} is not a heap name offset")
        need(keys[0] == 0, "first key of the leftmost node is the empty string")
        for i, child in enumerate(kids):
            if level > 0:
                node(child, level - 1)
                continue
            need(buf[child : child + 4] == b"SNOD", f"SNOD signature at {child:#x}")
            sver, srsv, nsym = u(buf, "BBH", child + 4)
            need(sver == 1 and srsv == 0 and 1 <= nsym <= 2 * leaf_k, f"symbol node version/count {sver}/{nsym}")
            need(child + 8 + 40 * 2 * leaf_k <= len(buf), "symbol node truncated (2K entries are allocated)")
            names = []
            for s in range(nsym):
                noff, haddr, ctype, ersv = u(buf, "QQII", child + 8 + 40 * s)
                need(noff < hsize and not in_free(noff) and ersv == 0, f"symbol entry {s}")
                need(ctype == 0, "dataset entries cache nothing")
                need(buf[child + 8 + 40 * s + 24 : child + 8 + 40 * s + 40] == b"\x00" * 16, "scratch pad")
                names.append(cstring(buf, hdata + noff))
                entries.append((names[-1], haddr))
            need(names == sorted(names), f"symbol node entries not sorted: {names}")
            lo, hi = cstring(buf, hdata + keys[i]), cstring(buf, hdata + keys[i + 1])
            need(all(lo < n <= hi for n in names) or (i == 0 and lo == "" and all(n <= hi for n in names)),
                 f"names {names} outside their key interval ({lo!r}, {hi!r}]")
            need(hi == names[-1], "right key is the largest name of the child (libhdf5 convention)")

    node(tree)
    need(len({n for n, _ in entries}) == len(entries), "duplicate link names")
    datasets = []
    spans = []
    for name, haddr in entries:
        d = parse_dataset(buf, haddr, name)
        daddr, dsize = d["data"]
        if dsize:
            spans.append((daddr, daddr + dsize, name))
        if d["dtype"]["class"] == 9:
            strings = []
            n = 1
            for s in d["shape"]:
                n *= s
            for i in range(n):
                length, gaddr, gidx = u(buf, "IQI", daddr + 16 * i)
                obj = parse_global_heap_object(buf, gaddr, gidx, f"dataset {name!r}")
                need(length <= len(obj), "vlen length exceeds the heap object")
                strings.append(obj[:length].decode("utf-8"))
            d["strings"] = strings
        datasets.append(d)
    spans.sort()
    for (a0, a1, n0), (b0, b1, n1) in zip(spans, spans[1:]):
        need(a1 <= b0, f"raw data of {n0!r} and {n1!r} overlap")
    return {"eof": eof, "datasets": datasets}


def signature(d):
    """What two files must agree on for one dataset, addresses aside."""
    return (d["name"], d["shape"], tuple(sorted((k, str(v)) for k, v in d["dtype"].items())),
            d["fill"], tuple(d["message_types"]), tuple(d["message_flags"]), d["data"][1])
