"""Independent STRUCTURAL validator for the HDF5 files this framework writes (Keras full-model checkpoints).

Neither h5py nor libhdf5 exists on the build or GPU machines, so nothing external can confirm that Keras would open our files.
This module is the next best thing: it is written from the HDF5 File Format Specification (version 0 superblock, version 1 object
headers, "old style" groups = symbol table message + local heap + version 1 B-tree + symbol-table nodes, contiguous layout,
version 1 attribute messages) and shares NO code with ``utils/hdf5.py`` or ``csrc/h5writer.cpp`` -- it re-derives every offset
from the bytes and checks the invariants libhdf5 relies on when it opens a file:

* superblock: signature, versions, 8-byte offsets/lengths, base address 0, end-of-file address == file size, root entry cached
  B-tree / heap addresses consistent with the root object header's symbol-table message;
* object headers: version 1, 8-byte aligned, message sizes multiples of 8, the sizes add up to the header's declared size, the
  declared message count matches, continuation blocks followed;
* groups: local heap signature / bounds, names NUL-terminated inside the heap data segment, B-tree node signature / type 0 / entry
  count <= 2K, keys in non-decreasing name order and bracketing their children, SNOD entries strictly sorted by name (libhdf5
  binary-searches them), every object reachable exactly once;
* datasets: dataspace + datatype + layout present, contiguous data inside the file, byte size == element count x element size;
* attributes: version 1 encoding, name / datatype / dataspace sizes padded to 8 bytes and consistent with the message size;
* no two allocated blocks overlap.

``validate(path)`` returns a report dict (counts) or raises ``H5FormatError`` naming the first violated rule.
"""
import struct

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEFINED = 0xFFFFFFFFFFFFFFFF

MSG_NIL, MSG_DATASPACE, MSG_DATATYPE, MSG_FILL_OLD, MSG_FILL, MSG_LAYOUT = 0x00, 0x01, 0x03, 0x04, 0x05, 0x08
MSG_ATTRIBUTE, MSG_CONTINUATION, MSG_SYMBOL_TABLE, MSG_MTIME = 0x0C, 0x10, 0x11, 0x12


class H5FormatError(ValueError):
    pass


class _Checker(object):
    def __init__(self, blob):
        self.b = blob
        self.size = len(blob)
        self.blocks = []               # (start, end, what) of everything that occupies file space
        self.report = {"groups": 0, "datasets": 0, "attributes": 0, "btree_nodes": 0, "snod_nodes": 0, "object_headers": 0,
                       "data_bytes": 0}
        self.seen_headers = set()

    # ------------------------------------------------------------------ primitives
    def need(self, cond, msg):
        if not cond:
            raise H5FormatError(msg)

    def u(self, off, n):
        self.need(0 <= off and off + n <= self.size, "read of %d bytes at %d runs past the end of the file (%d)" % (n, off, self.size))
        return int.from_bytes(self.b[off:off + n], "little")

    def claim(self, start, length, what):
        self.need(start >= 0 and start + length <= self.eof, "%s [%d, %d) lies outside the file (EOF address %d)" % (what, start, start + length, self.eof))
        if length:
            self.blocks.append((start, start + length, what))

    # ------------------------------------------------------------------ superblock
    def superblock(self):
        b = self.b
        self.need(b[:8] == SIGNATURE, "bad signature")
        self.need(b[8] == 0, "superblock version %d (expected 0)" % b[8])
        self.need(b[9] == 0 and b[10] == 0 and b[12] == 0, "free-space / root-entry / shared-header versions must be 0")
        self.need(b[13] == 8 and b[14] == 8, "size of offsets / lengths must be 8 (found %d / %d)" % (b[13], b[14]))
        self.leaf_k, self.internal_k = self.u(16, 2), self.u(18, 2)
        self.need(self.leaf_k > 0 and self.internal_k > 0, "group leaf / internal node K must be positive")
        base, free, eof, driver = self.u(24, 8), self.u(32, 8), self.u(40, 8), self.u(48, 8)
        self.need(base == 0, "base address must be 0")
        self.need(free == UNDEFINED, "free-space info address must be undefined")
        self.need(driver == UNDEFINED, "driver information address must be undefined")
        self.need(eof == self.size, "end-of-file address %d != file size %d" % (eof, self.size))
        self.eof = eof
        self.claim(0, 96, "superblock")
        # root group symbol table entry at 56: name offset (8), header address (8), cache type (4), reserved (4), scratch (16)
        root_header, cache = self.u(64, 8), self.u(72, 4)
        self.need(cache in (0, 1), "root entry cache type %d" % cache)
        cached = (self.u(80, 8), self.u(88, 8)) if cache == 1 else None
        return root_header, cached

    # ------------------------------------------------------------------ object headers
    def object_header(self, addr):
        """-> list of (type, flags, payload offset, payload size)"""
        self.need(addr % 8 == 0, "object header at %d is not 8-byte aligned" % addr)
        self.need(addr not in self.seen_headers, "object header at %d is linked twice (hard-link cycles are not written by this framework)" % addr)
        self.seen_headers.add(addr)
        self.report["object_headers"] += 1
        self.need(self.u(addr, 1) == 1, "object header at %d: version %d (expected 1)" % (addr, self.u(addr, 1)))
        self.need(self.u(addr + 1, 1) == 0, "object header at %d: reserved byte not zero" % addr)
        nmsg, refs, hsize = self.u(addr + 2, 2), self.u(addr + 4, 4), self.u(addr + 8, 4)
        self.need(refs >= 1, "object header at %d: reference count %d" % (addr, refs))
        self.claim(addr, 16 + hsize, "object header %d" % addr)
        msgs, chunks, count = [], [(addr + 16, hsize)], 0
        while chunks:
            off, left = chunks.pop(0)
            self.need(off % 8 == 0, "message block at %d is not 8-byte aligned" % off)
            while left > 0:
                self.need(left >= 8, "object header %d: %d stray bytes at the end of a message block" % (addr, left))
                mtype, msize, flags = self.u(off, 2), self.u(off + 2, 2), self.u(off + 4, 1)
                self.need(msize % 8 == 0, "object header %d: message type 0x%x has size %d (not a multiple of 8)" % (addr, mtype, msize))
                self.need(msize + 8 <= left, "object header %d: message type 0x%x (%d bytes) overruns its block" % (addr, mtype, msize))
                count += 1
                if mtype == MSG_CONTINUATION:
                    caddr, clen = self.u(off + 8, 8), self.u(off + 16, 8)
                    self.claim(caddr, clen, "continuation block of header %d" % addr)
                    chunks.append((caddr, clen))
                elif mtype != MSG_NIL:
                    msgs.append((mtype, flags, off + 8, msize))
                off += 8 + msize; left -= 8 + msize
        self.need(count == nmsg, "object header %d declares %d messages, %d found" % (addr, nmsg, count))
        return msgs

    # ------------------------------------------------------------------ groups
    def heap(self, addr):
        self.need(self.b[addr:addr + 4] == b"HEAP", "local heap at %d: bad signature" % addr)
        self.need(self.u(addr + 4, 1) == 0, "local heap at %d: version" % addr)
        dsize, free_head, daddr = self.u(addr + 8, 8), self.u(addr + 16, 8), self.u(addr + 24, 8)
        self.claim(addr, 32, "local heap header %d" % addr)
        self.claim(daddr, dsize, "local heap data of %d" % addr)
        self.need(free_head == UNDEFINED or free_head == 1 or free_head + 16 <= dsize, "local heap %d: free-list head %d outside the data segment" % (addr, free_head))
        return daddr, dsize

    def heap_name(self, heap, off):
        daddr, dsize = heap
        self.need(off < dsize, "name offset %d outside the heap data segment (%d bytes)" % (off, dsize))
        end = self.b.find(b"\x00", daddr + off, daddr + dsize)
        self.need(end >= 0, "name at heap offset %d is not NUL-terminated inside the heap" % off)
        return bytes(self.b[daddr + off:end])

    def btree(self, addr, heap, level_expected, out):
        """Walk a group B-tree node; appends (name, header address) in tree order; returns (first key name, last key name)."""
        self.need(self.b[addr:addr + 4] == b"TREE", "B-tree node at %d: bad signature" % addr)
        ntype, level, used = self.u(addr + 4, 1), self.u(addr + 5, 1), self.u(addr + 6, 2)
        self.need(ntype == 0, "B-tree node at %d: type %d (group nodes are type 0)" % (addr, ntype))
        self.need(level_expected is None or level == level_expected, "B-tree node at %d: level %d, parent expects %r" % (addr, level, level_expected))
        k = self.internal_k
        # an empty group is a root node (level 0, no parent) with zero entries and no symbol-table node; nowhere else may a node be empty
        self.need((1 if level_expected is not None or level > 0 else 0) <= used <= 2 * k, "B-tree node at %d: %d entries (K = %d)" % (addr, used, k))
        self.report["btree_nodes"] += 1
        self.claim(addr, 24 + (2 * k + 1) * 8 + 2 * k * 8, "B-tree node %d" % addr)
        keys = [self.heap_name(heap, self.u(addr + 24 + 16 * i, 8)) for i in range(used + 1)]
        kids = [self.u(addr + 24 + 16 * i + 8, 8) for i in range(used)]
        for i in range(used):
            self.need(keys[i] <= keys[i + 1], "B-tree node at %d: keys %r > %r out of order" % (addr, keys[i], keys[i + 1]))
            first = len(out)
            if level == 0:
                self.snod(kids[i], heap, out)
            else:
                self.btree(kids[i], heap, level - 1, out)
            names = [n for n, _ in out[first:]]
            self.need(names, "B-tree node at %d: child %d is empty" % (addr, i))
            # key[i] < every name in child i <= key[i+1]   (the very first key of a tree is the empty string)
            self.need(keys[i] < names[0] or (keys[i] == b"" and names[0] >= b""), "B-tree node at %d: child %d starts at %r, not above its left key %r" % (addr, i, names[0], keys[i]))
            self.need(names[-1] <= keys[i + 1], "B-tree node at %d: child %d ends at %r, above its right key %r" % (addr, i, names[-1], keys[i + 1]))
        return keys[0], keys[-1]

    def snod(self, addr, heap, out):
        self.need(self.b[addr:addr + 4] == b"SNOD", "symbol-table node at %d: bad signature" % addr)
        self.need(self.u(addr + 4, 1) == 1, "symbol-table node at %d: version" % addr)
        n = self.u(addr + 6, 2)
        self.need(1 <= n <= 2 * self.leaf_k, "symbol-table node at %d: %d symbols (leaf K = %d)" % (addr, n, self.leaf_k))
        self.report["snod_nodes"] += 1
        self.claim(addr, 8 + 2 * self.leaf_k * 40, "symbol-table node %d" % addr)
        prev = None
        for i in range(n):
            e = addr + 8 + 40 * i
            name = self.heap_name(heap, self.u(e, 8))
            self.need(name != b"", "symbol-table node at %d: empty link name" % addr)
            self.need(prev is None or prev < name, "symbol-table node at %d: names %r, %r not strictly increasing" % (addr, prev, name))
            prev = name
            cache = self.u(e + 16, 4)
            self.need(cache in (0, 1), "symbol-table node at %d: entry %r cache type %d" % (addr, name, cache))
            out.append((name, self.u(e + 8, 8)))

    # ------------------------------------------------------------------ messages
    def dataspace(self, off, size):
        ver, rank, flags = self.u(off, 1), self.u(off + 1, 1), self.u(off + 2, 1)
        self.need(ver == 1, "dataspace version %d" % ver)
        need = 8 + 8 * rank * (2 if flags & 1 else 1)
        self.need(need <= size, "dataspace message too short for rank %d" % rank)
        n = 1
        for i in range(rank):
            n *= self.u(off + 8 + 8 * i, 8)
        return n

    def datatype_size(self, off, size):
        self.need(size >= 8, "datatype message too short")
        cls_ver = self.u(off, 1)
        self.need(cls_ver >> 4 == 1, "datatype version %d" % (cls_ver >> 4))
        cls = cls_ver & 0x0F
        self.need(cls in (0, 1, 3), "datatype class %d is not fixed-point / float / string" % cls)
        esize = self.u(off + 4, 4)
        self.need(esize > 0, "datatype element size 0")
        if cls == 1:
            self.need(esize in (4, 8) and size >= 8 + 12, "floating-point datatype of %d bytes" % esize)
        return esize

    def attribute(self, off, size, owner):
        self.need(self.u(off, 1) == 1, "attribute of %s: version %d (expected 1)" % (owner, self.u(off, 1)))
        nlen, tlen, slen = self.u(off + 2, 2), self.u(off + 4, 2), self.u(off + 6, 2)
        pad = lambda v: (v + 7) & ~7
        self.need(nlen >= 2 and self.b[off + 8 + nlen - 1] == 0, "attribute of %s: name not NUL-terminated" % owner)
        t_off = off + 8 + pad(nlen); s_off = t_off + pad(tlen); d_off = s_off + pad(slen)
        self.need(d_off - off <= size, "attribute of %s: header parts (%d bytes) exceed the message (%d)" % (owner, d_off - off, size))
        esize = self.datatype_size(t_off, tlen)
        count = self.dataspace(s_off, slen)
        self.need(d_off - off + esize * count <= size, "attribute %r of %s: %d x %d data bytes do not fit in the message" %
                  (bytes(self.b[off + 8:off + 8 + nlen - 1]), owner, count, esize))
        self.report["attributes"] += 1

    # ------------------------------------------------------------------ objects
    def obj(self, addr, path, cached=None):
        msgs = self.object_header(addr)
        types = [m[0] for m in msgs]
        for mtype, flags, off, size in msgs:
            if mtype == MSG_ATTRIBUTE:
                self.attribute(off, size, path)
        if MSG_SYMBOL_TABLE in types:
            self.report["groups"] += 1
            _, _, off, size = msgs[types.index(MSG_SYMBOL_TABLE)]
            bt, hp = self.u(off, 8), self.u(off + 8, 8)
            if cached is not None:
                self.need(cached == (bt, hp), "root entry caches B-tree/heap %r but the root header says %r" % (cached, (bt, hp)))
            heap = self.heap(hp)
            links = []
            self.btree(bt, heap, None, links)
            names = [n for n, _ in links]
            self.need(names == sorted(names) and len(set(names)) == len(names), "group %s: links are not globally sorted / unique" % path)
            for name, child in links:
                self.obj(child, path.rstrip("/") + "/" + name.decode("utf8", "replace"))
        else:
            for must, label in ((MSG_DATASPACE, "dataspace"), (MSG_DATATYPE, "datatype"), (MSG_LAYOUT, "layout")):
                self.need(must in types, "dataset %s has no %s message" % (path, label))
            self.report["datasets"] += 1
            _, _, off, size = msgs[types.index(MSG_DATASPACE)]
            count = self.dataspace(off, size)
            _, _, off, size = msgs[types.index(MSG_DATATYPE)]
            esize = self.datatype_size(off, size)
            _, _, off, size = msgs[types.index(MSG_LAYOUT)]
            self.need(self.u(off, 1) == 3, "dataset %s: layout version %d (expected 3)" % (path, self.u(off, 1)))
            cls = self.u(off + 1, 1)
            self.need(cls == 1, "dataset %s: layout class %d (this framework writes contiguous = 1)" % (path, cls))
            daddr, dlen = self.u(off + 2, 8), self.u(off + 10, 8)
            self.need(dlen == count * esize, "dataset %s: %d data bytes for %d elements of %d bytes" % (path, dlen, count, esize))
            if dlen:
                self.need(daddr != UNDEFINED, "dataset %s: data address undefined for %d bytes" % (path, dlen))
                self.claim(daddr, dlen, "data of %s" % path)
                self.report["data_bytes"] += dlen

    def overlaps(self):
        blocks = sorted(self.blocks)
        for (s0, e0, w0), (s1, e1, w1) in zip(blocks, blocks[1:]):
            self.need(e0 <= s1, "%s [%d, %d) overlaps %s [%d, %d)" % (w0, s0, e0, w1, s1, e1))


def validate_bytes(blob):
    c = _Checker(bytes(blob))
    root, cached = c.superblock()
    c.obj(root, "/", cached)
    c.overlaps()
    c.report["file_bytes"] = c.size
    return c.report


def validate(path):
    with open(path, "rb") as f:
        return validate_bytes(f.read())


if __name__ == "__main__":
    import json
    import sys
    for p in sys.argv[1:]:
        print(p, json.dumps(validate(p)))
