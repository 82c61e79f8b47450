"""Minimal HDF5 writer/reader in pure Python + NumPy (h5py / libhdf5 are not available on the build machines).

Writer: the classic on-disk format that ``h5py.File(..., libver='earliest')`` / Keras 2.2 produce -- superblock v0,
version-1 object headers, symbol-table groups (local heap + v1 B-tree + SNOD), contiguous little-endian datasets,
attributes with fixed-length strings / numeric arrays.  That is exactly what a Keras full-model checkpoint needs
(reference rpv.py:101 ModelCheckpoint, DistHPO_mnist.ipynb:542 load_model; layout in SURVEY Appendix A).

Reader: the same, plus what real Keras / h5py files add: object-header continuation blocks, variable-length
strings (global heap), compact datasets, chunked datasets with deflate + shuffle filters (the ATLAS RPV files are
compressed), multi-level group B-trees, and version-2 object headers with compact (link-message) groups.

API is the h5py subset the framework uses: File(path, 'r'|'w'), group[name], create_group, create_dataset,
.attrs mapping, dataset[...] / .shape / .dtype, keys()/items()/__contains__, context manager.
"""
import struct
import zlib

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"
LEAF_K = 32          # symbol-table node holds up to 2*LEAF_K entries
INTERNAL_K = 16


def _pad8(n):
    return (n + 7) & ~7


# ======================================================================================================
# datatype / dataspace encoding
# ======================================================================================================
def _encode_dtype(dt):
    dt = np.dtype(dt)
    if dt.kind == "f":
        size = dt.itemsize
        if size == 4:
            props = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127); sign = 31
        elif size == 8:
            props = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023); sign = 63
        else:
            raise TypeError("unsupported float size")
        return struct.pack("<BBBBI", 0x11, 0x20, sign, 0, size) + props
    if dt.kind in "iu":
        bits = 0x08 if dt.kind == "i" else 0x00
        return struct.pack("<BBBBI", 0x10, bits, 0, 0, dt.itemsize) + struct.pack("<HH", 0, dt.itemsize * 8)
    if dt.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, max(dt.itemsize, 1))     # null-padded ASCII
    raise TypeError("unsupported dtype %r" % (dt,))


def _encode_space(shape):
    shape = tuple(int(s) for s in shape)
    return struct.pack("<BBBBI", 1, len(shape), 0, 0, 0) + b"".join(struct.pack("<Q", s) for s in shape)


def _as_array(value):
    if isinstance(value, str):
        value = value.encode("utf8")
    if isinstance(value, bytes):
        return np.array(value, dtype="S%d" % max(len(value), 1))
    arr = np.asarray(value)
    if arr.dtype.kind == "U":
        arr = np.char.encode(arr, "utf8")
    if arr.dtype.kind == "O":
        arr = np.array([v.encode("utf8") if isinstance(v, str) else v for v in arr.ravel()]).reshape(arr.shape)
    if arr.dtype == np.bool_:
        arr = arr.astype(np.uint8)
    if arr.dtype.kind in "fiu" and arr.dtype.byteorder == ">":
        arr = arr.astype(arr.dtype.newbyteorder("<"))
    return arr


# ======================================================================================================
# writer
# ======================================================================================================
class _WAttrs(dict):
    def __setitem__(self, k, v):
        dict.__setitem__(self, k, _as_array(v))


class _WNode(object):
    def __init__(self):
        self.attrs = _WAttrs()


class _WDataset(_WNode):
    def __init__(self, data):
        _WNode.__init__(self)
        self.data = np.ascontiguousarray(_as_array(data))
        self.shape, self.dtype = self.data.shape, self.data.dtype


class _WGroup(_WNode):
    def __init__(self):
        _WNode.__init__(self)
        self.children = {}

    def _walk(self, path, create):
        node = self
        for part in [p for p in path.split("/") if p]:
            if part not in node.children:
                if not create:
                    raise KeyError(path)
                node.children[part] = _WGroup()
            node = node.children[part]
        return node

    def create_group(self, name):
        return self._walk(name, True)

    def require_group(self, name):
        return self._walk(name, True)

    def create_dataset(self, name, data=None, shape=None, dtype=None):
        parts = [p for p in name.split("/") if p]
        parent = self._walk("/".join(parts[:-1]), True)
        if data is None:
            data = np.zeros(shape, dtype or np.float32)
        elif dtype is not None:
            data = np.asarray(data, dtype)
        ds = _WDataset(data)
        parent.children[parts[-1]] = ds
        return ds

    def __getitem__(self, name):
        return self._walk(name, False)

    def __contains__(self, name):
        try:
            self._walk(name, False); return True
        except KeyError:
            return False

    def keys(self):
        return list(self.children.keys())


class _Writer(object):
    def __init__(self):
        self.buf = bytearray(96)      # superblock placeholder

    def alloc(self, data):
        while len(self.buf) % 8:
            self.buf.append(0)
        addr = len(self.buf)
        self.buf += data
        return addr

    @staticmethod
    def _msg(mtype, body, flags=0):
        body = bytes(body) + b"\0" * (_pad8(len(body)) - len(body))
        if len(body) > 0xFFF8:
            raise ValueError("HDF5 object-header message too large (%d bytes); split the attribute" % len(body))
        return struct.pack("<HHBBBB", mtype, len(body), flags, 0, 0, 0) + body

    def _attr_msgs(self, attrs):
        out = []
        for name, arr in attrs.items():
            nm = name.encode("utf8") + b"\0"
            dt = _encode_dtype(arr.dtype)
            sp = _encode_space(arr.shape)
            body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(sp))
            body += nm + b"\0" * (_pad8(len(nm)) - len(nm))
            body += dt + b"\0" * (_pad8(len(dt)) - len(dt))
            body += sp + b"\0" * (_pad8(len(sp)) - len(sp))
            body += np.ascontiguousarray(arr).tobytes()
            out.append(self._msg(0x000C, body))
        return out

    def _header(self, msgs):
        data = b"".join(msgs)
        return struct.pack("<BBHII", 1, 0, len(msgs), 1, len(data)) + b"\0\0\0\0" + data

    def write_dataset(self, ds):
        raw = ds.data.tobytes()
        daddr = self.alloc(raw) if raw else UNDEF
        msgs = [self._msg(0x0001, _encode_space(ds.shape)),
                self._msg(0x0003, _encode_dtype(ds.dtype), flags=1),
                self._msg(0x0005, struct.pack("<BBBB", 2, 2, 2, 0)),
                self._msg(0x0008, struct.pack("<BBQQ", 3, 1, daddr, len(raw)))]
        msgs += self._attr_msgs(ds.attrs)
        return self.alloc(self._header(msgs))

    def write_group(self, grp):
        """Returns (object header address, btree address, heap address)."""
        entries = []
        for name in sorted(grp.children, key=lambda s: s.encode("utf8")):
            child = grp.children[name]
            if isinstance(child, _WGroup):
                oh, bt, hp = self.write_group(child)
                entries.append((name, oh, 1, bt, hp))
            else:
                entries.append((name, self.write_dataset(child), 0, 0, 0))
        # local heap data: offset 0 = empty string
        heap = bytearray(8)
        offs = []
        for name, _, _, _, _ in entries:
            offs.append(len(heap))
            nm = name.encode("utf8") + b"\0"
            heap += nm + b"\0" * (_pad8(len(nm)) - len(nm))
        heap_data_addr = self.alloc(bytes(heap))
        heap_addr = self.alloc(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, len(heap), 1, heap_data_addr))
        # symbol-table nodes
        per = 2 * LEAF_K
        # an empty group is a B-tree root with zero entries and no symbol-table node at all (what libhdf5 itself creates)
        chunks = [list(range(i, min(i + per, len(entries)))) for i in range(0, len(entries), per)]
        if len(chunks) > 2 * INTERNAL_K:
            raise ValueError("too many entries in one HDF5 group for this writer")
        snods, keys = [], [0]
        for ch in chunks:
            body = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(ch)))
            for i in ch:
                name, oh, ctype, bt, hp = entries[i]
                scratch = struct.pack("<QQ", bt, hp) if ctype == 1 else b"\0" * 16
                body += struct.pack("<QQII", offs[i], oh, ctype, 0) + scratch
            body += b"\0" * (40 * (per - len(ch)))
            snods.append(self.alloc(bytes(body)))
            keys.append(offs[ch[-1]] if ch else 0)
        node = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, 0, len(snods), UNDEF, UNDEF))
        for i, a in enumerate(snods):
            node += struct.pack("<QQ", keys[i], a)
        node += struct.pack("<Q", keys[len(snods)])
        node += b"\0" * (24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8 - len(node))
        bt_addr = self.alloc(bytes(node))
        msgs = [self._msg(0x0011, struct.pack("<QQ", bt_addr, heap_addr))] + self._attr_msgs(grp.attrs)
        return self.alloc(self._header(msgs)), bt_addr, heap_addr

    def finish(self, root):
        oh, bt, hp = self.write_group(root)
        while len(self.buf) % 8:
            self.buf.append(0)
        sb = SIG + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, LEAF_K, INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack("<QQII", 0, oh, 1, 0) + struct.pack("<QQ", bt, hp)
        assert len(sb) == 96
        self.buf[0:96] = sb
        return bytes(self.buf)


# ======================================================================================================
# reader
# ======================================================================================================
class _RFile(object):
    def __init__(self, path):
        self.fh = open(path, "rb")
        head = self.read(0, 64)
        if head[:8] != SIG:
            raise IOError("%s is not an HDF5 file" % path)
        ver = head[8]
        if ver in (0, 1):
            self.so, self.sl = head[13], head[14]
            off = 24 + (4 if ver == 1 else 0)
            sb = self.read(0, off + 4 * self.so + 2 * self.so + 8 + 16 + 16)
            self.base = self._u(sb, off, self.so)
            root_entry = off + 4 * self.so
            self.root_addr = self._u(sb, root_entry + self.so, self.so)
        elif ver in (2, 3):
            self.so, self.sl = head[9], head[10]
            sb = self.read(0, 12 + 4 * self.so + 4)
            self.base = self._u(sb, 12, self.so)
            self.root_addr = self._u(sb, 12 + 3 * self.so, self.so)
        else:
            raise IOError("unsupported HDF5 superblock version %d" % ver)
        if self.so != 8 or self.sl != 8:
            raise IOError("only 8-byte offsets/lengths are supported")

    @staticmethod
    def _u(b, off, n):
        return int.from_bytes(b[off:off + n], "little")

    def read(self, addr, n):
        self.fh.seek(addr)
        return self.fh.read(n)

    def close(self):
        self.fh.close()

    # ---------------------------------------------------------------- object headers
    def messages(self, addr):
        """List of (type, flags, body-bytes) for the object header at ``addr`` (v1 or v2)."""
        head = self.read(addr, 16)
        out = []
        if head[:4] == b"OHDR":
            return self._messages_v2(addr)
        nmsgs = struct.unpack_from("<H", head, 2)[0]
        size = struct.unpack_from("<I", head, 8)[0]
        blocks = [(addr + 16, size)]
        while blocks and len(out) < nmsgs + 64:
            baddr, bsize = blocks.pop(0)
            data = self.read(baddr, bsize)
            p = 0
            while p + 8 <= len(data):
                mtype, msize, flags = struct.unpack_from("<HHB", data, p)
                body = data[p + 8:p + 8 + msize]
                p += 8 + msize
                if mtype == 0x0010:      # continuation
                    caddr, clen = struct.unpack_from("<QQ", body, 0)
                    blocks.append((caddr, clen))
                else:
                    out.append((mtype, flags, body))
        return out

    def _messages_v2(self, addr):
        head = self.read(addr, 64)
        flags = head[5]
        p = 6
        if flags & 0x20:
            p += 16
        if flags & 0x10:
            p += 4
        szbytes = 1 << (flags & 3)
        chunk0 = self._u(head, p, szbytes); p += szbytes
        track_order = bool(flags & 0x04)
        out = []
        blocks = [(addr + p, chunk0)]
        while blocks:
            baddr, bsize = blocks.pop(0)
            data = self.read(baddr, bsize)
            q = 0
            while q + 4 <= len(data):
                mtype = data[q]; msize = struct.unpack_from("<H", data, q + 1)[0]; mflags = data[q + 3]
                q += 4 + (2 if track_order else 0)
                body = data[q:q + msize]; q += msize
                if mtype == 0x10:
                    caddr, clen = struct.unpack_from("<QQ", body, 0)
                    blocks.append((caddr + 4, clen - 8))      # skip "OCHK", drop checksum
                elif mtype != 0:
                    out.append((mtype, mflags, body))
        return out

    # ---------------------------------------------------------------- datatype / dataspace
    def parse_dtype(self, b):
        cls = b[0] & 0x0F
        bits0 = b[1]
        size = struct.unpack_from("<I", b, 4)[0]
        if cls == 0:
            order = ">" if bits0 & 1 else "<"
            return np.dtype("%s%s%d" % (order, "i" if bits0 & 0x08 else "u", size)), None
        if cls == 1:
            order = ">" if bits0 & 1 else "<"
            return np.dtype("%sf%d" % (order, size)), None
        if cls == 3:
            return np.dtype("S%d" % size), None
        if cls == 9:
            vtype = bits0 & 0x0F
            if vtype == 1:
                return np.dtype("O"), "vlen_str"
            base, _ = self.parse_dtype(b[8:])
            return np.dtype("O"), ("vlen", base)
        if cls == 8:    # enum (h5py stores bool as enum of int8)
            base, _ = self.parse_dtype(b[8:])
            return base, None
        raise TypeError("unsupported HDF5 datatype class %d" % cls)

    @staticmethod
    def parse_space(b):
        ver = b[0]
        rank = b[1]
        if ver == 1:
            off = 8
        else:
            off = 4
            if b[3] == 2 if ver == 2 else False:
                return None
        return tuple(struct.unpack_from("<Q", b, off + 8 * i)[0] for i in range(rank))

    def global_heap_object(self, coll_addr, index):
        head = self.read(coll_addr, 16)
        if head[:4] != b"GCOL":
            raise IOError("bad global heap collection")
        size = struct.unpack_from("<Q", head, 8)[0]
        data = self.read(coll_addr, size)
        p = 16
        while p + 16 <= len(data):
            idx, _, _, osize = struct.unpack_from("<HHIQ", data, p)
            if idx == 0:
                break
            if idx == index:
                return data[p + 16:p + 16 + osize]
            p += 16 + _pad8(osize)
        raise KeyError("global heap object %d not found" % index)

    def decode_values(self, raw, dtype, special, shape):
        n = int(np.prod(shape)) if shape else 1
        if special is None:
            arr = np.frombuffer(raw, dtype=dtype, count=n).reshape(shape)
            if arr.dtype.byteorder == ">":
                arr = arr.astype(arr.dtype.newbyteorder("<"))
            return arr.copy()
        vals = []
        for i in range(n):
            length, gaddr, gidx = struct.unpack_from("<IQI", raw, 16 * i)
            blob = self.global_heap_object(gaddr, gidx)[:length] if gaddr not in (0, UNDEF) else b""
            if special == "vlen_str":
                vals.append(bytes(blob))
            else:
                vals.append(np.frombuffer(blob, dtype=special[1]).copy())
        if not shape:
            return vals[0]
        out = np.empty(n, dtype=object)
        for i, v in enumerate(vals):
            out[i] = v
        return out.reshape(shape)

    # ---------------------------------------------------------------- groups
    def group_links(self, msgs):
        links = {}
        for mtype, _, body in msgs:
            if mtype == 0x0011:
                bt, hp = struct.unpack_from("<QQ", body, 0)
                heap = self.read(hp, 32)
                dsize, _, daddr = struct.unpack_from("<QQQ", heap, 8)
                hdata = self.read(daddr, dsize)
                self._walk_btree(bt, hdata, links)
            elif mtype == 0x0006:       # link message (compact new-style groups)
                p = 0
                ver, flags = body[0], body[1]; p = 2
                ltype = 0
                if flags & 0x08:
                    ltype = body[p]; p += 1
                if flags & 0x04:
                    p += 8
                if flags & 0x10:
                    p += 1
                lsz = 1 << (flags & 3)
                nlen = self._u(body, p, lsz); p += lsz
                name = body[p:p + nlen].decode("utf8"); p += nlen
                if ltype == 0:
                    links[name] = self._u(body, p, 8)
            elif mtype == 0x0002:
                raise IOError("dense (fractal-heap) groups are not supported by this reader")
        return links

    def _walk_btree(self, addr, hdata, links):
        node = self.read(addr, 24)
        if node[:4] != b"TREE":
            raise IOError("bad group B-tree node")
        level, used = node[5], struct.unpack_from("<H", node, 6)[0]
        body = self.read(addr + 24, (2 * used + 1) * 8)
        for i in range(used):
            child = struct.unpack_from("<Q", body, 8 + 16 * i)[0]
            if level > 0:
                self._walk_btree(child, hdata, links)
            else:
                sn = self.read(child, 8)
                nsym = struct.unpack_from("<H", sn, 6)[0]
                ent = self.read(child + 8, 40 * nsym)
                for j in range(nsym):
                    noff, oh = struct.unpack_from("<QQ", ent, 40 * j)
                    end = hdata.index(b"\0", noff)
                    links[hdata[noff:end].decode("utf8")] = oh

    def attributes(self, msgs):
        out = {}
        for mtype, _, body in msgs:
            if mtype != 0x000C:
                continue
            ver = body[0]
            nsz, dsz, ssz = struct.unpack_from("<HHH", body, 2)
            p = 8
            if ver == 3:
                p = 9
            pad = _pad8 if ver == 1 else (lambda v: v)
            name = body[p:p + nsz].split(b"\0")[0].decode("utf8"); p += pad(nsz)
            dt, special = self.parse_dtype(body[p:p + dsz]); p += pad(dsz)
            shape = self.parse_space(body[p:p + ssz]); p += pad(ssz)
            if shape is None:
                out[name] = None
                continue
            val = self.decode_values(body[p:], dt, special, shape)
            out[name] = val[()] if isinstance(val, np.ndarray) and val.shape == () else val
        return out


class _RNode(object):
    def __init__(self, rf, addr, name):
        self._rf, self._addr, self.name = rf, addr, name
        self._msgs = rf.messages(addr)
        self._attrs = None

    @property
    def attrs(self):
        if self._attrs is None:
            self._attrs = self._rf.attributes(self._msgs)
        return self._attrs


class Group(_RNode):
    def __init__(self, rf, addr, name):
        _RNode.__init__(self, rf, addr, name)
        self._links = rf.group_links(self._msgs)

    def keys(self):
        return sorted(self._links.keys())

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        return len(self._links)

    def __contains__(self, name):
        try:
            self[name]; return True
        except KeyError:
            return False

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def __getitem__(self, path):
        node = self
        for part in [p for p in path.split("/") if p]:
            if not isinstance(node, Group) or part not in node._links:
                raise KeyError(path)
            addr = node._links[part]
            msgs = node._rf.messages(addr)
            is_group = any(m[0] in (0x0011, 0x0002, 0x0006, 0x000A) for m in msgs) and not any(m[0] == 0x0008 for m in msgs)
            child_name = (node.name.rstrip("/") + "/" + part)
            node = Group(node._rf, addr, child_name) if is_group else Dataset(node._rf, addr, child_name)
        return node


class Dataset(_RNode):
    def __init__(self, rf, addr, name):
        _RNode.__init__(self, rf, addr, name)
        self._filters = []
        for mtype, _, body in self._msgs:
            if mtype == 0x0001:
                self.shape = rf.parse_space(body)
            elif mtype == 0x0003:
                self.dtype, self._special = rf.parse_dtype(body)
            elif mtype == 0x0008:
                self._layout = body
            elif mtype == 0x000B:
                self._filters = self._parse_filters(body)

    @staticmethod
    def _parse_filters(b):
        ver, n = b[0], b[1]
        p = 8 if ver == 1 else 2
        out = []
        for _ in range(n):
            fid = struct.unpack_from("<H", b, p)[0]; p += 2
            nlen = 0
            if ver == 1 or fid >= 256:
                nlen = struct.unpack_from("<H", b, p)[0]; p += 2
            flags, ncv = struct.unpack_from("<HH", b, p); p += 4
            if nlen:
                p += _pad8(nlen) if ver == 1 else nlen
            cvals = struct.unpack_from("<%dI" % ncv, b, p); p += 4 * ncv
            if ver == 1 and ncv % 2:
                p += 4
            out.append((fid, cvals))
        return out

    def __len__(self):
        return self.shape[0]

    def _read_all(self):
        rf = self._rf
        lay = self._layout
        ver = lay[0]
        n = int(np.prod(self.shape)) if self.shape else 1
        item = 16 if self._special else self.dtype.itemsize
        if ver == 3:
            cls = lay[1]
            if cls == 1:
                addr, size = struct.unpack_from("<QQ", lay, 2)
                raw = rf.read(addr, n * item) if addr != UNDEF else b"\0" * (n * item)
            elif cls == 0:
                size = struct.unpack_from("<H", lay, 2)[0]
                raw = lay[4:4 + size]
            elif cls == 2:
                rank = lay[2]
                bt = struct.unpack_from("<Q", lay, 3)[0]
                cdims = struct.unpack_from("<%dI" % rank, lay, 11)
                return self._read_chunked(bt, cdims[:-1])
            else:
                raise IOError("unsupported layout class %d" % cls)
        elif ver in (1, 2):
            rank, cls = lay[1], lay[2]
            p = 8
            addr = UNDEF
            if cls != 0:
                addr = struct.unpack_from("<Q", lay, p)[0]; p += 8
            dims = struct.unpack_from("<%dI" % rank, lay, p)
            if cls == 2:
                return self._read_chunked(addr, dims[:-1])
            raw = rf.read(addr, n * item)
        else:
            raise IOError("unsupported data layout version %d" % ver)
        return rf.decode_values(raw, self.dtype, self._special, self.shape)

    def _read_chunked(self, btree, cdims):
        out = np.zeros(self.shape, dtype=self.dtype.newbyteorder("<") if self.dtype.kind in "fiu" else self.dtype)
        if btree == UNDEF:
            return out
        rank = len(self.shape)
        self._walk_chunks(btree, rank, cdims, out)
        return out

    def _walk_chunks(self, addr, rank, cdims, out):
        rf = self._rf
        node = rf.read(addr, 24)
        if node[:4] != b"TREE":
            raise IOError("bad chunk B-tree node")
        level, used = node[5], struct.unpack_from("<H", node, 6)[0]
        ksz = 8 + 8 * (rank + 1)
        body = rf.read(addr + 24, used * (ksz + 8) + ksz)
        for i in range(used):
            k = i * (ksz + 8)
            csize, fmask = struct.unpack_from("<II", body, k)
            offs = struct.unpack_from("<%dQ" % rank, body, k + 8)
            child = struct.unpack_from("<Q", body, k + ksz)[0]
            if level > 0:
                self._walk_chunks(child, rank, cdims, out)
                continue
            raw = rf.read(child, csize)
            for fi, (fid, cvals) in reversed(list(enumerate(self._filters))):
                if fmask & (1 << fi):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    es = cvals[0] if cvals else self.dtype.itemsize
                    a = np.frombuffer(raw, np.uint8)
                    nel = len(a) // es
                    raw = a[:nel * es].reshape(es, nel).T.tobytes() + a[nel * es:].tobytes()
                elif fid == 3:
                    raw = raw[:-4]
                else:
                    raise IOError("unsupported HDF5 filter id %d" % fid)
            chunk = np.frombuffer(raw, dtype=self.dtype, count=int(np.prod(cdims))).reshape(cdims)
            sl_out = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, cdims, self.shape))
            sl_in = tuple(slice(0, s.stop - s.start) for s in sl_out)
            out[sl_out] = chunk[sl_in]

    def _contiguous_addr(self):
        lay = self._layout
        if lay[0] == 3 and lay[1] == 1 and not self._special:
            return struct.unpack_from("<Q", lay, 2)[0]
        return None

    def __getitem__(self, key):
        # fast path: leading-axis slice of a contiguous dataset reads only the bytes it needs (load_file(n_samples))
        if isinstance(key, slice) and self.shape and key.step in (None, 1):
            addr = self._contiguous_addr()
            if addr is not None and addr != UNDEF:
                start, stop, _ = key.indices(self.shape[0])
                stop = max(stop, start)
                row = int(np.prod(self.shape[1:])) if len(self.shape) > 1 else 1
                raw = self._rf.read(addr + start * row * self.dtype.itemsize, (stop - start) * row * self.dtype.itemsize)
                return self._rf.decode_values(raw, self.dtype, None, (stop - start,) + tuple(self.shape[1:]))
        full = self._read_all()
        if isinstance(key, tuple) and key == ():
            return full[()] if isinstance(full, np.ndarray) and full.shape == () else full
        if key is Ellipsis:
            return full
        return full[key]

    @property
    def value(self):
        return self[()]


# ======================================================================================================
# native writer binding (csrc/h5writer.cpp -> lib/libcibh5.so).  CIB_H5_NATIVE=0 forces the Python writer.
# ======================================================================================================
_native = [None]


def native_lib():
    """The C++ writer library, or None when it is not built."""
    import ctypes as C
    import os
    if _native[0] is None:
        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libcibh5.so")
        if os.environ.get("CIB_H5_NATIVE", "1") == "0" or not os.path.exists(path):
            _native[0] = False
        else:
            L = C.CDLL(path)
            L.cibh5_new.restype = C.c_void_p
            L.cibh5_error.restype = C.c_char_p
            L.cibh5_free.argtypes = [C.c_void_p]
            L.cibh5_group.argtypes = [C.c_void_p, C.c_char_p]
            L.cibh5_dataset.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_longlong), C.c_void_p]
            L.cibh5_attr.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_longlong), C.c_void_p]
            L.cibh5_serialize.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong]; L.cibh5_serialize.restype = C.c_longlong
            L.cibh5_write_file.argtypes = [C.c_void_p, C.c_char_p]
            if hasattr(L, "cibh5r_open"):                   # reader half (csrc/h5reader.cpp)
                ll = C.POINTER(C.c_longlong); pi = C.POINTER(C.c_int)
                L.cibh5r_open.restype = C.c_void_p; L.cibh5r_open.argtypes = [C.c_char_p]
                L.cibh5r_error.restype = C.c_char_p
                L.cibh5r_close.argtypes = [C.c_void_p]
                L.cibh5r_count.argtypes = [C.c_void_p]
                L.cibh5r_object.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), pi, pi, pi, pi, ll, ll, ll]
                L.cibh5r_attr.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_char_p), pi, pi, pi, ll, ll, ll]
                L.cibh5r_read.argtypes = [C.c_void_p, C.c_longlong, C.c_longlong, C.c_void_p]; L.cibh5r_read.restype = C.c_longlong
            _native[0] = L
    return _native[0] or None


def _native_tree(L, root):
    """Mirror a _WGroup tree into the C++ builder; returns the handle (caller frees) or None if something is unsupported."""
    import ctypes as C
    t = L.cibh5_new()

    def arr_args(a):
        a = np.require(a, requirements="C")          # (ascontiguousarray would turn 0-d arrays into 1-d)
        kind = {"f": 0, "i": 1, "u": 2, "S": 3}.get(a.dtype.kind)
        if kind is None:
            raise TypeError(a.dtype)
        shape = (C.c_longlong * max(a.ndim, 1))(*a.shape)
        return kind, a.dtype.itemsize, a.ndim, shape, a.ctypes.data_as(C.c_void_p), a

    def attrs(path, node):
        for k, v in node.attrs.items():
            kind, isz, nd, shape, ptr, keep = arr_args(v)
            if L.cibh5_attr(t, path.encode("utf8"), k.encode("utf8"), kind, isz, nd, shape, ptr):
                raise RuntimeError(L.cibh5_error().decode())

    def visit(path, grp):
        attrs(path, grp)
        for name, child in grp.children.items():
            cp = path.rstrip("/") + "/" + name
            if isinstance(child, _WGroup):
                if L.cibh5_group(t, cp.encode("utf8")):
                    raise RuntimeError(L.cibh5_error().decode())
                visit(cp, child)
            else:
                kind, isz, nd, shape, ptr, keep = arr_args(child.data)
                if L.cibh5_dataset(t, cp.encode("utf8"), kind, isz, nd, shape, ptr):
                    raise RuntimeError(L.cibh5_error().decode())
                attrs(cp, child)
    try:
        visit("/", root)
    except (TypeError, RuntimeError):
        L.cibh5_free(t)
        return None
    return t


def _native_write(root, path):
    L = native_lib()
    if L is None:
        return False
    t = _native_tree(L, root)
    if t is None:
        return False
    try:
        return L.cibh5_write_file(t, path.encode("utf8")) == 0
    finally:
        L.cibh5_free(t)


def native_serialize(root):
    """Bytes produced by the C++ writer for a _WGroup tree (tests compare them with _Writer().finish(root))."""
    import ctypes as C
    L = native_lib()
    t = _native_tree(L, root)
    try:
        n = L.cibh5_serialize(t, None, 0)
        buf = (C.c_ubyte * n)()
        L.cibh5_serialize(t, buf, n)
        return bytes(buf)
    finally:
        L.cibh5_free(t)


# ======================================================================================================
# native reader (csrc/h5reader.cpp): the C++ side indexes the file, these classes map the bytes
# ======================================================================================================
_KIND_DTYPE = {0: "f", 1: "i", 2: "u"}


class _NIndex(object):
    """One pass of the C++ reader over a classic-format file: every object's path, dtype, shape, data offset and attributes."""

    def __init__(self, L, path):
        import ctypes as C
        self.L = L
        self.h = L.cibh5r_open(path.encode("utf8"))
        if not self.h:
            raise IOError(L.cibh5r_error().decode("utf8", "replace"))
        self.children, self.nodes = {}, {}
        name = C.c_char_p(); isg, kind, isz, rank = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        shape = (C.c_longlong * 8)(); off, ln = C.c_longlong(), C.c_longlong()
        for i in range(L.cibh5r_count(self.h)):
            na = L.cibh5r_object(self.h, i, C.byref(name), C.byref(isg), C.byref(kind), C.byref(isz), C.byref(rank), shape, C.byref(off), C.byref(ln))
            p = name.value.decode("utf8")
            attrs = []
            aname = C.c_char_p(); ak, ai, ar = C.c_int(), C.c_int(), C.c_int(); ash = (C.c_longlong * 8)(); ao, al = C.c_longlong(), C.c_longlong()
            for j in range(na):
                L.cibh5r_attr(self.h, i, j, C.byref(aname), C.byref(ak), C.byref(ai), C.byref(ar), ash, C.byref(ao), C.byref(al))
                attrs.append((aname.value.decode("utf8"), ak.value, ai.value, tuple(ash[d] for d in range(ar.value)), ao.value, al.value))
            self.nodes[p] = (bool(isg.value), kind.value, isz.value, tuple(shape[d] for d in range(rank.value)), off.value, ln.value, attrs)
            if p != "/":
                parent, leaf = p.rsplit("/", 1)
                self.children.setdefault(parent or "/", []).append(leaf)
            self.children.setdefault(p, []) if isg.value else None

    @staticmethod
    def dtype_of(kind, itemsize):
        return np.dtype("S%d" % itemsize) if kind == 3 else np.dtype("<%s%d" % (_KIND_DTYPE[kind], itemsize))

    def read(self, off, ln):
        import ctypes as C
        buf = (C.c_ubyte * max(ln, 1))()
        got = self.L.cibh5r_read(self.h, off, ln, buf)
        if got != ln:
            raise IOError("short read from the HDF5 file")
        return bytes(buf[:ln])

    def values(self, kind, itemsize, shape, off, ln):
        arr = np.frombuffer(self.read(off, ln), dtype=self.dtype_of(kind, itemsize)).reshape(shape).copy()
        return arr

    def close(self):
        if self.h:
            self.L.cibh5r_close(self.h); self.h = None


class _NNode(object):
    def __init__(self, idx, path):
        self._idx, self.name = idx, path
        self._attrs = None

    @property
    def attrs(self):
        if self._attrs is None:
            out = {}
            for name, kind, isz, shape, off, ln in self._idx.nodes[self.name][6]:
                v = self._idx.values(kind, isz, shape, off, ln)
                out[name] = v[()] if v.shape == () else v
            self._attrs = out
        return self._attrs


class _NGroup(_NNode):
    def keys(self):
        return sorted(self._idx.children.get(self.name, []))

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        return len(self._idx.children.get(self.name, []))

    def __contains__(self, name):
        try:
            self[name]; return True
        except KeyError:
            return False

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def __getitem__(self, path):
        full = self.name.rstrip("/")
        for part in [p for p in path.split("/") if p]:
            full = full + "/" + part
        if full == "":
            full = "/"
        node = self._idx.nodes.get(full)
        if node is None:
            raise KeyError(path)
        return _NGroup(self._idx, full) if node[0] else _NDataset(self._idx, full)


class _NDataset(_NNode):
    def __init__(self, idx, path):
        _NNode.__init__(self, idx, path)
        _, kind, isz, shape, off, ln, _ = idx.nodes[path]
        self.shape, self.dtype = shape, idx.dtype_of(kind, isz)
        self._kind, self._isz, self._off, self._len = kind, isz, off, ln

    def __len__(self):
        return self.shape[0] if self.shape else 0

    def _read_all(self):
        return self._idx.values(self._kind, self._isz, self.shape, self._off, self._len)

    def __getitem__(self, key):
        # leading-axis slices of a contiguous dataset read only the rows they need
        if isinstance(key, slice) and self.shape and key.step in (None, 1):
            lo, hi, _ = key.indices(self.shape[0])
            row = int(np.prod(self.shape[1:], dtype=np.int64)) * self._isz
            return self._idx.values(self._kind, self._isz, (max(hi - lo, 0),) + tuple(self.shape[1:]), self._off + lo * row, max(hi - lo, 0) * row)
        return self._read_all()[key]

    @property
    def value(self):
        return self._read_all()


class File(object):
    """``File(path, 'r')`` -> read-only view; ``File(path, 'w')`` -> in-memory tree serialised on close().

    Reading goes through the native C++ index (csrc/h5reader.cpp) when the file stays inside the classic subset it understands
    (every checkpoint this framework or Keras 2.2 writes); anything else -- chunked / compressed datasets, version-2 object headers,
    variable-length strings -- is handled by the pure-Python reader below.  ``CIB_H5_NATIVE=0`` forces the Python reader."""

    def __init__(self, path, mode="r"):
        self.path, self.mode = path, mode
        self._rf = self._nidx = None
        if mode == "r":
            L = native_lib()
            if L is not None and hasattr(L, "cibh5r_open"):
                try:
                    self._nidx = _NIndex(L, path)
                    self._root = _NGroup(self._nidx, "/")
                except IOError as e:
                    if not str(e).startswith("unsupported:"):
                        # not an HDF5 file / damaged: let the Python reader produce its own (more detailed) error
                        pass
                    self._nidx = None
            if self._nidx is None:
                self._rf = _RFile(path)
                self._root = Group(self._rf, self._rf.root_addr + self._rf.base, "/")
        elif mode == "w":
            self._root = _WGroup()
        else:
            raise ValueError("mode must be 'r' or 'w'")
        self._closed = False

    def __getattr__(self, item):
        return getattr(self._root, item)

    def __getitem__(self, k):
        return self._root[k]

    def __contains__(self, k):
        return k in self._root

    def __iter__(self):
        return iter(self._root.keys())

    def flush(self):
        pass

    def close(self):
        if self._closed:
            return
        self._closed = True
        if self.mode == "w":
            if _native_write(self._root, self.path):       # C++ writer (csrc/h5writer.cpp), same bytes
                return
            data = _Writer().finish(self._root)
            tmp = self.path + ".tmp~"
            with open(tmp, "wb") as fh:
                fh.write(data)
            import os
            os.replace(tmp, self.path)
        elif self._nidx is not None:
            self._nidx.close()
        else:
            self._rf.close()

    @property
    def reader(self):
        """'native' or 'python' (read mode): which implementation serves this file."""
        return "native" if self._nidx is not None else "python" if self._rf is not None else None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
