// Native HDF5 writer (host only, no libhdf5): the classic on-disk format that Keras 2.2 / h5py(libver='earliest') produce --
// superblock v0, version-1 object headers, symbol-table groups (local heap + v1 B-tree + SNOD), contiguous little-endian
// datasets, attributes with fixed-length strings / numeric arrays.  It writes the Keras full-model checkpoints
// (ModelCheckpoint every epoch, reference rpv.py:100-101, DistHPO_mnist.ipynb:184-185; layout in SURVEY Appendix A) and is
// byte-for-byte identical to the pure-Python writer in utils/hdf5.py, which stays as the fall-back and as the test oracle.
//
// C ABI (cibh5_*): build a tree of groups / datasets / attributes, then serialise it to a file (written to <name>.tmp and
// renamed, so a reader never sees a half-written checkpoint).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace {

constexpr uint64_t UNDEF = 0xFFFFFFFFFFFFFFFFull;
constexpr int LEAF_K = 32, INTERNAL_K = 16;
thread_local std::string g_err;

using Bytes = std::vector<uint8_t>;
inline size_t pad8(size_t n) { return (n + 7) & ~size_t(7); }
template <class T> void put(Bytes& b, T v) { const uint8_t* p = reinterpret_cast<const uint8_t*>(&v); b.insert(b.end(), p, p + sizeof(T)); }
inline void put_bytes(Bytes& b, const void* p, size_t n) { const uint8_t* q = static_cast<const uint8_t*>(p); b.insert(b.end(), q, q + n); }
inline void pad_to8(Bytes& b, size_t from) { b.resize(from + pad8(b.size() - from), 0); }

struct Array {
  int kind = 0;            // 0 float, 1 signed int, 2 unsigned int, 3 fixed-length string
  int itemsize = 4;
  std::vector<uint64_t> shape;
  Bytes data;
};

struct Node {
  bool is_group = true;
  Array arr;                                     // datasets
  std::vector<std::pair<std::string, Array>> attrs;   // insertion order, like the Python dict
  std::map<std::string, std::unique_ptr<Node>> children;   // std::map orders by bytes, as the B-tree needs
};

Bytes encode_dtype(const Array& a) {
  Bytes b;
  if (a.kind == 0) {
    const bool f32 = a.itemsize == 4;
    put<uint8_t>(b, 0x11); put<uint8_t>(b, 0x20); put<uint8_t>(b, f32 ? 31 : 63); put<uint8_t>(b, 0); put<uint32_t>(b, a.itemsize);
    put<uint16_t>(b, 0); put<uint16_t>(b, f32 ? 32 : 64); put<uint8_t>(b, f32 ? 23 : 52); put<uint8_t>(b, f32 ? 8 : 11);
    put<uint8_t>(b, 0); put<uint8_t>(b, f32 ? 23 : 52); put<uint32_t>(b, f32 ? 127 : 1023);
  } else if (a.kind == 1 || a.kind == 2) {
    put<uint8_t>(b, 0x10); put<uint8_t>(b, a.kind == 1 ? 0x08 : 0x00); put<uint8_t>(b, 0); put<uint8_t>(b, 0); put<uint32_t>(b, a.itemsize);
    put<uint16_t>(b, 0); put<uint16_t>(b, (uint16_t)(a.itemsize * 8));
  } else {
    put<uint8_t>(b, 0x13); put<uint8_t>(b, 0x01); put<uint8_t>(b, 0); put<uint8_t>(b, 0); put<uint32_t>(b, a.itemsize > 0 ? a.itemsize : 1);
  }
  return b;
}

Bytes encode_space(const std::vector<uint64_t>& shape) {
  Bytes b;
  put<uint8_t>(b, 1); put<uint8_t>(b, (uint8_t)shape.size()); put<uint8_t>(b, 0); put<uint8_t>(b, 0); put<uint32_t>(b, 0);
  for (uint64_t s : shape) put<uint64_t>(b, s);
  return b;
}

struct Writer {
  Bytes buf = Bytes(96, 0);      // superblock placeholder
  bool failed = false;

  uint64_t alloc(const Bytes& data) {
    buf.resize(pad8(buf.size()), 0);
    const uint64_t addr = buf.size();
    buf.insert(buf.end(), data.begin(), data.end());
    return addr;
  }
  Bytes msg(uint16_t type, const Bytes& body_in, uint8_t flags = 0) {
    Bytes body = body_in; body.resize(pad8(body.size()), 0);
    if (body.size() > 0xFFF8) { failed = true; g_err = "HDF5 object-header message too large (" + std::to_string(body.size()) + " bytes); split the attribute"; }
    Bytes m; put<uint16_t>(m, type); put<uint16_t>(m, (uint16_t)body.size()); put<uint8_t>(m, flags); put<uint8_t>(m, 0); put<uint8_t>(m, 0); put<uint8_t>(m, 0);
    m.insert(m.end(), body.begin(), body.end());
    return m;
  }
  std::vector<Bytes> attr_msgs(const Node& n) {
    std::vector<Bytes> out;
    for (const auto& kv : n.attrs) {
      Bytes nm(kv.first.begin(), kv.first.end()); nm.push_back(0);
      const Bytes dt = encode_dtype(kv.second), sp = encode_space(kv.second.shape);
      Bytes body; put<uint8_t>(body, 1); put<uint8_t>(body, 0); put<uint16_t>(body, (uint16_t)nm.size()); put<uint16_t>(body, (uint16_t)dt.size()); put<uint16_t>(body, (uint16_t)sp.size());
      size_t at = body.size(); body.insert(body.end(), nm.begin(), nm.end()); pad_to8(body, at);
      at = body.size(); body.insert(body.end(), dt.begin(), dt.end()); pad_to8(body, at);
      at = body.size(); body.insert(body.end(), sp.begin(), sp.end()); pad_to8(body, at);
      body.insert(body.end(), kv.second.data.begin(), kv.second.data.end());
      out.push_back(msg(0x000C, body));
    }
    return out;
  }
  Bytes header(const std::vector<Bytes>& msgs) {
    size_t total = 0; for (const auto& m : msgs) total += m.size();
    Bytes h; put<uint8_t>(h, 1); put<uint8_t>(h, 0); put<uint16_t>(h, (uint16_t)msgs.size()); put<uint32_t>(h, 1); put<uint32_t>(h, (uint32_t)total);
    put<uint32_t>(h, 0);
    for (const auto& m : msgs) h.insert(h.end(), m.begin(), m.end());
    return h;
  }
  uint64_t write_dataset(const Node& ds) {
    const uint64_t daddr = ds.arr.data.empty() ? UNDEF : alloc(ds.arr.data);
    std::vector<Bytes> msgs;
    msgs.push_back(msg(0x0001, encode_space(ds.arr.shape)));
    msgs.push_back(msg(0x0003, encode_dtype(ds.arr), 1));
    { Bytes b; put<uint8_t>(b, 2); put<uint8_t>(b, 2); put<uint8_t>(b, 2); put<uint8_t>(b, 0); msgs.push_back(msg(0x0005, b)); }
    { Bytes b; put<uint8_t>(b, 3); put<uint8_t>(b, 1); put<uint64_t>(b, daddr); put<uint64_t>(b, (uint64_t)ds.arr.data.size()); msgs.push_back(msg(0x0008, b)); }
    for (auto& m : attr_msgs(ds)) msgs.push_back(std::move(m));
    return alloc(header(msgs));
  }
  struct Entry { std::string name; uint64_t oh; uint32_t ctype; uint64_t bt, hp; };
  void write_group(const Node& g, uint64_t* oh_out, uint64_t* bt_out, uint64_t* hp_out) {
    std::vector<Entry> entries;
    for (const auto& kv : g.children) {
      if (kv.second->is_group) { uint64_t oh, bt, hp; write_group(*kv.second, &oh, &bt, &hp); entries.push_back({kv.first, oh, 1u, bt, hp}); }
      else entries.push_back({kv.first, write_dataset(*kv.second), 0u, 0, 0});
    }
    Bytes heap(8, 0);                              // local heap data: offset 0 = empty string
    std::vector<uint64_t> offs;
    for (const auto& e : entries) {
      offs.push_back(heap.size());
      const size_t at = heap.size();
      heap.insert(heap.end(), e.name.begin(), e.name.end()); heap.push_back(0);
      pad_to8(heap, at);
    }
    const uint64_t heap_data = alloc(heap);
    Bytes hh; put_bytes(hh, "HEAP", 4); put<uint8_t>(hh, 0); put<uint8_t>(hh, 0); put<uint8_t>(hh, 0); put<uint8_t>(hh, 0);
    put<uint64_t>(hh, (uint64_t)heap.size()); put<uint64_t>(hh, 1); put<uint64_t>(hh, heap_data);
    const uint64_t heap_addr = alloc(hh);
    const size_t per = 2 * LEAF_K;
    std::vector<std::vector<size_t>> chunks;
    for (size_t i = 0; i < entries.size(); i += per) { std::vector<size_t> c; for (size_t j = i; j < entries.size() && j < i + per; ++j) c.push_back(j); chunks.push_back(c); }
    // an empty group is a B-tree root with zero entries and no symbol-table node at all (what libhdf5 itself creates)
    if (chunks.size() > 2 * INTERNAL_K) { failed = true; g_err = "too many entries in one HDF5 group for this writer"; }
    std::vector<uint64_t> snods, keys{0};
    for (const auto& ch : chunks) {
      Bytes body; put_bytes(body, "SNOD", 4); put<uint8_t>(body, 1); put<uint8_t>(body, 0); put<uint16_t>(body, (uint16_t)ch.size());
      for (size_t i : ch) {
        const Entry& e = entries[i];
        put<uint64_t>(body, offs[i]); put<uint64_t>(body, e.oh); put<uint32_t>(body, e.ctype); put<uint32_t>(body, 0);
        if (e.ctype == 1) { put<uint64_t>(body, e.bt); put<uint64_t>(body, e.hp); } else { put<uint64_t>(body, 0); put<uint64_t>(body, 0); }
      }
      body.resize(body.size() + 40 * (per - ch.size()), 0);
      snods.push_back(alloc(body));
      keys.push_back(ch.empty() ? 0 : offs[ch.back()]);
    }
    Bytes node; put_bytes(node, "TREE", 4); put<uint8_t>(node, 0); put<uint8_t>(node, 0); put<uint16_t>(node, (uint16_t)snods.size()); put<uint64_t>(node, UNDEF); put<uint64_t>(node, UNDEF);
    for (size_t i = 0; i < snods.size(); ++i) { put<uint64_t>(node, keys[i]); put<uint64_t>(node, snods[i]); }
    put<uint64_t>(node, keys[snods.size()]);
    node.resize(24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8, 0);
    const uint64_t bt_addr = alloc(node);
    std::vector<Bytes> msgs;
    { Bytes b; put<uint64_t>(b, bt_addr); put<uint64_t>(b, heap_addr); msgs.push_back(msg(0x0011, b)); }
    for (auto& m : attr_msgs(g)) msgs.push_back(std::move(m));
    *oh_out = alloc(header(msgs)); *bt_out = bt_addr; *hp_out = heap_addr;
  }
  void finish(const Node& root) {
    uint64_t oh, bt, hp; write_group(root, &oh, &bt, &hp);
    buf.resize(pad8(buf.size()), 0);
    Bytes sb; put_bytes(sb, "\x89HDF\r\n\x1a\n", 8);
    const uint8_t v[8] = {0, 0, 0, 0, 0, 8, 8, 0}; put_bytes(sb, v, 8);
    put<uint16_t>(sb, LEAF_K); put<uint16_t>(sb, INTERNAL_K); put<uint32_t>(sb, 0);
    put<uint64_t>(sb, 0); put<uint64_t>(sb, UNDEF); put<uint64_t>(sb, (uint64_t)buf.size()); put<uint64_t>(sb, UNDEF);
    put<uint64_t>(sb, 0); put<uint64_t>(sb, oh); put<uint32_t>(sb, 1); put<uint32_t>(sb, 0); put<uint64_t>(sb, bt); put<uint64_t>(sb, hp);
    memcpy(buf.data(), sb.data(), 96);
  }
};

struct Tree { Node root; };

Node* walk(Node& root, const char* path, bool create, bool make_dataset_leaf) {
  Node* node = &root;
  std::string p(path ? path : "");
  size_t i = 0;
  std::vector<std::string> parts;
  while (i < p.size()) { size_t j = p.find('/', i); if (j == std::string::npos) j = p.size(); if (j > i) parts.push_back(p.substr(i, j - i)); i = j + 1; }
  for (size_t k = 0; k < parts.size(); ++k) {
    if (!node->is_group) return nullptr;
    auto it = node->children.find(parts[k]);
    if (it == node->children.end()) {
      if (!create) return nullptr;
      auto nn = std::make_unique<Node>();
      nn->is_group = !(make_dataset_leaf && k + 1 == parts.size());
      it = node->children.emplace(parts[k], std::move(nn)).first;
    } else if (make_dataset_leaf && k + 1 == parts.size()) {
      it->second = std::make_unique<Node>(); it->second->is_group = false;       // replace, like the Python writer
    }
    node = it->second.get();
  }
  return node;
}

bool fill(Array& a, int kind, int itemsize, int rank, const long long* shape, const void* data) {
  if (kind < 0 || kind > 3 || itemsize < 0 || rank < 0 || rank > 8) { g_err = "bad array description"; return false; }
  if ((kind == 0 && itemsize != 4 && itemsize != 8) || ((kind == 1 || kind == 2) && itemsize != 1 && itemsize != 2 && itemsize != 4 && itemsize != 8)) { g_err = "unsupported item size"; return false; }
  a.kind = kind; a.itemsize = kind == 3 && itemsize < 1 ? 1 : itemsize;
  a.shape.assign(shape, shape + rank);
  uint64_t n = 1; for (int i = 0; i < rank; ++i) n *= (uint64_t)shape[i];
  const size_t bytes = (size_t)n * (size_t)a.itemsize;
  a.data.assign(static_cast<const uint8_t*>(data), static_cast<const uint8_t*>(data) + (data ? bytes : 0));
  if (!data) a.data.assign(bytes, 0);
  return true;
}

}  // namespace

extern "C" {

const char* cibh5_error(void) { return g_err.c_str(); }
void* cibh5_new(void) { return new Tree(); }
void cibh5_free(void* t) { delete static_cast<Tree*>(t); }

int cibh5_group(void* t, const char* path) {
  Node* n = walk(static_cast<Tree*>(t)->root, path, true, false);
  if (!n || !n->is_group) { g_err = std::string("cannot create group ") + path; return 1; }
  return 0;
}
int cibh5_dataset(void* t, const char* path, int kind, int itemsize, int rank, const long long* shape, const void* data) {
  Node* n = walk(static_cast<Tree*>(t)->root, path, true, true);
  if (!n || n->is_group) { g_err = std::string("cannot create dataset ") + path; return 1; }
  return fill(n->arr, kind, itemsize, rank, shape, data) ? 0 : 1;
}
int cibh5_attr(void* t, const char* node_path, const char* name, int kind, int itemsize, int rank, const long long* shape, const void* data) {
  Node* n = walk(static_cast<Tree*>(t)->root, node_path, false, false);
  if (!n) { g_err = std::string("no such node ") + node_path; return 1; }
  Array a; if (!fill(a, kind, itemsize, rank, shape, data)) return 1;
  for (auto& kv : n->attrs) if (kv.first == name) { kv.second = std::move(a); return 0; }
  n->attrs.emplace_back(name, std::move(a));
  return 0;
}
// Serialise; returns the size and, when out != nullptr, copies at most cap bytes (used by the tests to compare with the
// Python writer).  cibh5_write_file writes <path>.tmp and renames it over <path>.
long long cibh5_serialize(void* t, void* out, long long cap) {
  Writer w; w.finish(static_cast<Tree*>(t)->root);
  if (w.failed) return -1;
  if (out) memcpy(out, w.buf.data(), (size_t)std::min<long long>(cap, (long long)w.buf.size()));
  return (long long)w.buf.size();
}
int cibh5_write_file(void* t, const char* path) {
  Writer w; w.finish(static_cast<Tree*>(t)->root);
  if (w.failed) return 1;
  const std::string tmp = std::string(path) + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) { g_err = "cannot open " + tmp; return 1; }
  const size_t n = fwrite(w.buf.data(), 1, w.buf.size(), f);
  if (fclose(f) != 0 || n != w.buf.size()) { g_err = "short write to " + tmp; remove(tmp.c_str()); return 1; }
  if (rename(tmp.c_str(), path) != 0) { g_err = "cannot rename " + tmp; return 1; }
  return 0;
}

}  // extern "C"
