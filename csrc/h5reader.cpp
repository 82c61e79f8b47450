// Native HDF5 reader (host only, no libhdf5) for the classic on-disk format -- what this framework's writers, Keras 2.2 and
// h5py(libver='earliest') produce for a full-model checkpoint (reference DistHPO_mnist.ipynb:542 load_model, rpv.py:101
// ModelCheckpoint; layout in SURVEY Appendix A): superblock v0/v1, version-1 object headers with continuation blocks,
// symbol-table groups (local heap + multi-level v1 B-tree + symbol-table nodes), little-endian fixed-point / floating-point /
// fixed-length string datatypes, simple dataspaces, contiguous and compact layouts, version-1 attributes.
//
// It builds an INDEX of the file in one pass (object paths, dtypes, shapes, file offsets of the raw data and of every attribute
// value); the Python side (utils/hdf5.py) maps the bytes from those offsets.  Anything outside the subset -- version-2 object
// headers, chunked / filtered datasets (the compressed ATLAS files), variable-length strings, big-endian data, shared messages --
// makes cibh5r_open fail with a message starting "unsupported:", and the pure-Python reader takes over.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

constexpr uint64_t UNDEF = 0xFFFFFFFFFFFFFFFFull;
thread_local std::string g_rerr;

struct RAttr {
  std::string name;
  int kind = 0, itemsize = 0;                 // kind: 0 float, 1 signed int, 2 unsigned int, 3 fixed-length string
  std::vector<uint64_t> shape;
  uint64_t off = 0, len = 0;                  // raw value bytes in the file
};
struct RObj {
  std::string path;
  bool group = true;
  int kind = 0, itemsize = 0;
  std::vector<uint64_t> shape;
  uint64_t off = 0, len = 0;                  // contiguous / compact data
  std::vector<RAttr> attrs;
};

struct Reader {
  FILE* f = nullptr;
  uint64_t size = 0;
  std::vector<RObj> objs;
  int depth = 0;

  bool fail(const std::string& m) { g_rerr = m; return false; }
  bool read(uint64_t off, size_t n, std::vector<uint8_t>& out) {
    if (off == UNDEF || off + n > size) return fail("read past the end of the file");
    out.resize(n);
    if (fseeko(f, (off_t)off, SEEK_SET) != 0 || (n && fread(out.data(), 1, n, f) != n)) return fail("short read");
    return true;
  }
  static uint64_t u(const uint8_t* p, int n) { uint64_t v = 0; for (int i = n - 1; i >= 0; --i) v = (v << 8) | p[i]; return v; }

  struct Msg { int type; int flags; uint64_t off; size_t len; };      // payload position in the file

  bool messages(uint64_t addr, std::vector<Msg>& out) {
    std::vector<uint8_t> h;
    if (!read(addr, 16, h)) return false;
    if (h[0] == 'O' && h[1] == 'H' && h[2] == 'D' && h[3] == 'R') return fail("unsupported: version-2 object header");
    if (h[0] != 1) return fail("unsupported: object header version " + std::to_string(h[0]));
    const int nmsg = (int)u(&h[2], 2);
    std::vector<std::pair<uint64_t, uint64_t>> chunks{{addr + 16, u(&h[8], 4)}};
    int seen = 0;
    for (size_t c = 0; c < chunks.size(); ++c) {
      std::vector<uint8_t> b;
      if (!read(chunks[c].first, (size_t)chunks[c].second, b)) return false;
      size_t p = 0;
      while (p + 8 <= b.size() && seen < nmsg) {
        const int type = (int)u(&b[p], 2); const size_t len = (size_t)u(&b[p + 2], 2); const int flags = b[p + 4];
        if (p + 8 + len > b.size()) return fail("object header message overruns its block");
        ++seen;
        if (flags & 0x02) return fail("unsupported: shared object header message");
        if (type == 0x0010) chunks.push_back({u(&b[p + 8], 8), u(&b[p + 16], 8)});
        else if (type != 0) out.push_back({type, flags, chunks[c].first + p + 8, len});
        p += 8 + len;
      }
    }
    return true;
  }

  bool dtype(const uint8_t* b, size_t n, int* kind, int* itemsize) {
    if (n < 8) return fail("datatype message too short");
    const int cls = b[0] & 0x0F, bits0 = b[1];
    *itemsize = (int)u(b + 4, 4);
    if (cls == 0) { if (bits0 & 1) return fail("unsupported: big-endian integers"); *kind = (bits0 & 0x08) ? 1 : 2; return true; }
    if (cls == 1) { if (bits0 & 1) return fail("unsupported: big-endian floats"); *kind = 0; return true; }
    if (cls == 3) { *kind = 3; return true; }
    return fail("unsupported: datatype class " + std::to_string(cls));
  }
  bool space(const uint8_t* b, size_t n, std::vector<uint64_t>* shape) {
    if (n < 4) return fail("dataspace message too short");
    const int ver = b[0], rank = b[1];
    size_t off = 8;
    if (ver == 2) { off = 4; if (b[3] == 2) return fail("unsupported: null dataspace"); }
    else if (ver != 1) return fail("unsupported: dataspace version " + std::to_string(ver));
    if (off + 8 * (size_t)rank > n) return fail("dataspace message too short for its rank");
    shape->clear();
    for (int i = 0; i < rank; ++i) shape->push_back(u(b + off + 8 * i, 8));
    return true;
  }
  static uint64_t count(const std::vector<uint64_t>& s) { uint64_t n = 1; for (uint64_t v : s) n *= v; return n; }

  bool attribute(const Msg& m, RAttr* a) {
    std::vector<uint8_t> b;
    if (!read(m.off, m.len, b)) return false;
    if (b.size() < 8 || b[0] != 1) return fail("unsupported: attribute message version " + std::to_string(b.empty() ? -1 : b[0]));
    const size_t nsz = (size_t)u(&b[2], 2), dsz = (size_t)u(&b[4], 2), ssz = (size_t)u(&b[6], 2);
    auto pad8 = [](size_t v) { return (v + 7) & ~size_t(7); };
    size_t p = 8;
    if (p + pad8(nsz) + pad8(dsz) + pad8(ssz) > b.size()) return fail("attribute message too short");
    a->name.assign(reinterpret_cast<const char*>(&b[p]), strnlen(reinterpret_cast<const char*>(&b[p]), nsz)); p += pad8(nsz);
    if (!dtype(&b[p], dsz, &a->kind, &a->itemsize)) return false; p += pad8(dsz);
    if (!space(&b[p], ssz, &a->shape)) return false; p += pad8(ssz);
    a->len = count(a->shape) * (uint64_t)a->itemsize;
    if (p + a->len > b.size()) return fail("attribute value overruns its message");
    a->off = m.off + p;
    return true;
  }

  bool heap_name(const std::vector<uint8_t>& heap, uint64_t off, std::string* out) {
    if (off >= heap.size()) return fail("link name outside the local heap");
    out->assign(reinterpret_cast<const char*>(&heap[off]), strnlen(reinterpret_cast<const char*>(&heap[off]), heap.size() - off));
    return true;
  }
  bool btree(uint64_t addr, const std::vector<uint8_t>& heap, std::vector<std::pair<std::string, uint64_t>>* links, int level_guard) {
    if (level_guard > 16) return fail("group B-tree deeper than 16 levels");
    std::vector<uint8_t> h;
    if (!read(addr, 24, h)) return false;
    if (memcmp(h.data(), "TREE", 4) != 0 || h[4] != 0) return fail("bad group B-tree node");
    const int level = h[5], used = (int)u(&h[6], 2);
    std::vector<uint8_t> body;
    if (!read(addr + 24, (size_t)(2 * used + 1) * 8, body)) return false;
    for (int i = 0; i < used; ++i) {
      const uint64_t child = u(&body[8 + 16 * i], 8);
      if (level > 0) { if (!btree(child, heap, links, level_guard + 1)) return false; continue; }
      std::vector<uint8_t> sn;
      if (!read(child, 8, sn)) return false;
      if (memcmp(sn.data(), "SNOD", 4) != 0) return fail("bad symbol-table node");
      const int nsym = (int)u(&sn[6], 2);
      std::vector<uint8_t> ent;
      if (!read(child + 8, (size_t)nsym * 40, ent)) return false;
      for (int j = 0; j < nsym; ++j) {
        std::string name;
        if (!heap_name(heap, u(&ent[40 * j], 8), &name)) return false;
        links->push_back({name, u(&ent[40 * j + 8], 8)});
      }
    }
    return true;
  }

  bool object(uint64_t addr, const std::string& path) {
    if (++depth > 64) return fail("group nesting deeper than 64");
    std::vector<Msg> msgs;
    if (!messages(addr, msgs)) return false;
    RObj o; o.path = path;
    const Msg* stab = nullptr; const Msg* lay = nullptr; const Msg* dt = nullptr; const Msg* sp = nullptr;
    for (const Msg& m : msgs) {
      if (m.type == 0x0011) stab = &m;
      else if (m.type == 0x0008) lay = &m;
      else if (m.type == 0x0003) dt = &m;
      else if (m.type == 0x0001) sp = &m;
      else if (m.type == 0x000B) return fail("unsupported: filtered dataset " + path);
      else if (m.type == 0x0002 || m.type == 0x0006 || m.type == 0x000A) return fail("unsupported: new-style group " + path);
      else if (m.type == 0x000C) { RAttr a; if (!attribute(m, &a)) return false; o.attrs.push_back(std::move(a)); }
    }
    if (stab && !lay) {
      o.group = true;
      const size_t self = objs.size();
      objs.push_back(std::move(o));
      std::vector<uint8_t> b;
      if (!read(stab->off, 16, b)) return false;
      const uint64_t bt = u(&b[0], 8), hp = u(&b[8], 8);
      std::vector<uint8_t> hh;
      if (!read(hp, 32, hh)) return false;
      if (memcmp(hh.data(), "HEAP", 4) != 0) return fail("bad local heap");
      std::vector<uint8_t> heap;
      if (!read(u(&hh[24], 8), (size_t)u(&hh[8], 8), heap)) return false;
      std::vector<std::pair<std::string, uint64_t>> links;
      if (!btree(bt, heap, &links, 0)) return false;
      (void)self;
      for (auto& kv : links)
        if (!object(kv.second, (path == "/" ? std::string("/") : path + "/") + kv.first)) return false;
    } else {
      if (!lay || !dt || !sp) return fail("object " + path + " is neither a symbol-table group nor a complete dataset");
      o.group = false;
      std::vector<uint8_t> b;
      if (!read(dt->off, dt->len, b)) return false;
      if (!dtype(b.data(), b.size(), &o.kind, &o.itemsize)) return false;
      if (!read(sp->off, sp->len, b)) return false;
      if (!space(b.data(), b.size(), &o.shape)) return false;
      if (!read(lay->off, lay->len, b)) return false;
      if (b.size() < 2 || b[0] != 3) return fail("unsupported: data layout message version " + std::to_string(b.empty() ? -1 : b[0]));
      if (b[1] == 1) { o.off = u(&b[2], 8); o.len = u(&b[10], 8); }
      else if (b[1] == 0) { o.len = u(&b[2], 2); o.off = lay->off + 4; }
      else return fail("unsupported: chunked dataset " + path);
      if (o.len != count(o.shape) * (uint64_t)o.itemsize) return fail("dataset " + path + ": data size does not match its shape");
      if (o.len && (o.off == UNDEF || o.off + o.len > size)) return fail("dataset " + path + ": data outside the file");
      objs.push_back(std::move(o));
    }
    --depth;
    return true;
  }

  bool open(const char* path) {
    f = fopen(path, "rb");
    if (!f) return fail(std::string("cannot open ") + path);
    fseeko(f, 0, SEEK_END); size = (uint64_t)ftello(f);
    std::vector<uint8_t> sb;
    if (!read(0, 96, sb)) return false;
    static const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    if (memcmp(sb.data(), sig, 8) != 0) return fail("not an HDF5 file (signature at offset 0)");
    const int ver = sb[8];
    if (ver > 1) return fail("unsupported: superblock version " + std::to_string(ver));
    if (sb[13] != 8 || sb[14] != 8) return fail("unsupported: offsets / lengths are not 8 bytes");
    const size_t base = ver == 0 ? 24 : 28;           // v1 has 4 more bytes (indexed storage K + reserved)
    const uint64_t root = u(&sb[base + 32 + 8], 8);   // root symbol-table entry: name offset (8), object header address (8)
    return object(root, "/");
  }
  ~Reader() { if (f) fclose(f); }
};

}  // namespace

extern "C" {

const char* cibh5r_error(void) { return g_rerr.c_str(); }

void* cibh5r_open(const char* path) {
  Reader* r = new Reader();
  if (!r->open(path)) { delete r; return nullptr; }
  return r;
}
void cibh5r_close(void* h) { delete static_cast<Reader*>(h); }
int cibh5r_count(void* h) { return (int)static_cast<Reader*>(h)->objs.size(); }

// shape8 receives up to 8 extents.  Returns the number of attributes, or -1 for a bad index.
int cibh5r_object(void* h, int i, const char** path, int* is_group, int* kind, int* itemsize, int* rank, long long* shape8,
                  long long* off, long long* len) {
  Reader* r = static_cast<Reader*>(h);
  if (i < 0 || i >= (int)r->objs.size()) return -1;
  const RObj& o = r->objs[i];
  *path = o.path.c_str(); *is_group = o.group ? 1 : 0; *kind = o.kind; *itemsize = o.itemsize; *rank = (int)o.shape.size();
  for (size_t d = 0; d < o.shape.size() && d < 8; ++d) shape8[d] = (long long)o.shape[d];
  *off = (long long)o.off; *len = (long long)o.len;
  return (int)o.attrs.size();
}
int cibh5r_attr(void* h, int i, int j, const char** name, int* kind, int* itemsize, int* rank, long long* shape8, long long* off, long long* len) {
  Reader* r = static_cast<Reader*>(h);
  if (i < 0 || i >= (int)r->objs.size() || j < 0 || j >= (int)r->objs[i].attrs.size()) return 1;
  const RAttr& a = r->objs[i].attrs[j];
  *name = a.name.c_str(); *kind = a.kind; *itemsize = a.itemsize; *rank = (int)a.shape.size();
  for (size_t d = 0; d < a.shape.size() && d < 8; ++d) shape8[d] = (long long)a.shape[d];
  *off = (long long)a.off; *len = (long long)a.len;
  return 0;
}
// Raw bytes [off, off + len) of the file into out; returns the number of bytes read or -1.
long long cibh5r_read(void* h, long long off, long long len, void* out) {
  Reader* r = static_cast<Reader*>(h);
  if (off < 0 || len < 0 || (uint64_t)(off + len) > r->size) return -1;
  if (fseeko(r->f, (off_t)off, SEEK_SET) != 0) return -1;
  return (long long)fread(out, 1, (size_t)len, r->f);
}

}  // extern "C"
