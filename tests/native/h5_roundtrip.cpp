// Host-side native code under AddressSanitizer / UndefinedBehaviorSanitizer: builds a checkpoint-shaped tree with the C++ HDF5 writer
// (csrc/h5writer.cpp), writes it, indexes it again with the C++ reader (csrc/h5reader.cpp) and compares every dataset and attribute.
// Compiled and run by tests/test_native_sanitizers.py with -fsanitize=address,undefined: any out-of-bounds access, leak or undefined
// shift in the (de)serialisers aborts the run.  (sanitizers for the native runtime.)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

extern "C" {
const char* cibh5_error(void);
void* cibh5_new(void);
void cibh5_free(void*);
int cibh5_group(void*, const char*);
int cibh5_dataset(void*, const char*, int, int, int, const long long*, const void*);
int cibh5_attr(void*, const char*, const char*, int, int, int, const long long*, const void*);
int cibh5_write_file(void*, const char*);
const char* cibh5r_error(void);
void* cibh5r_open(const char*);
void cibh5r_close(void*);
int cibh5r_count(void*);
int cibh5r_object(void*, int, const char**, int*, int*, int*, int*, long long*, long long*, long long*);
int cibh5r_attr(void*, int, int, const char**, int*, int*, int*, long long*, long long*, long long*);
long long cibh5r_read(void*, long long, long long, void*);
}

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "FAILED %s:%d: %s  [%s | %s]\n", __FILE__, __LINE__, #c, cibh5_error(), cibh5r_error()); return 1; } } while (0)

int main(int argc, char** argv) {
  const std::string path = argc > 1 ? argv[1] : "/tmp/cib_h5_roundtrip.h5";
  void* t = cibh5_new();
  CHECK(t);
  // a group with more links than one symbol-table node holds, an empty group, nested groups, several dtypes
  CHECK(cibh5_group(t, "/model_weights") == 0);
  CHECK(cibh5_group(t, "/model_weights/dropout_1") == 0);
  std::vector<std::vector<float>> payloads;
  for (int i = 0; i < 150; ++i) {
    std::vector<float> v((size_t)i + 1);
    for (size_t k = 0; k < v.size(); ++k) v[k] = (float)(i * 1000 + (int)k);
    payloads.push_back(v);
    char name[64]; snprintf(name, sizeof name, "/model_weights/w%03d", i);
    const long long shape[1] = {(long long)v.size()};
    CHECK(cibh5_dataset(t, name, 0, 4, 1, shape, v.data()) == 0);
  }
  const long long s2[2] = {3, 5};
  std::vector<int64_t> ints(15); for (int k = 0; k < 15; ++k) ints[k] = 7 - k;
  CHECK(cibh5_dataset(t, "/optimizer_weights/iterations", 1, 8, 2, s2, ints.data()) == 0);
  const char* cfg = "{\"class_name\": \"Model\", \"config\": {}}";
  const long long s0[1] = {0};
  CHECK(cibh5_attr(t, "/", "model_config", 3, (int)strlen(cfg), 0, s0, cfg) == 0);
  const char names[3][8] = {"conv_1", "pool_1", "dense_1"};
  const long long s3[1] = {3};
  CHECK(cibh5_attr(t, "/model_weights", "layer_names", 3, 8, 1, s3, names) == 0);
  const double lr = 0.001;
  CHECK(cibh5_attr(t, "/optimizer_weights/iterations", "lr", 0, 8, 0, s0, &lr) == 0);
  CHECK(cibh5_write_file(t, path.c_str()) == 0);
  cibh5_free(t);

  void* r = cibh5r_open(path.c_str());
  CHECK(r);
  int datasets = 0, groups = 0, attrs = 0;
  for (int i = 0; i < cibh5r_count(r); ++i) {
    const char* p; int isg, kind, isz, rank; long long shape[8], off, len;
    const int na = cibh5r_object(r, i, &p, &isg, &kind, &isz, &rank, shape, &off, &len);
    CHECK(na >= 0);
    attrs += na;
    for (int j = 0; j < na; ++j) {
      const char* an; int ak, ai, ar; long long ash[8], ao, al;
      CHECK(cibh5r_attr(r, i, j, &an, &ak, &ai, &ar, ash, &ao, &al) == 0);
      std::vector<char> buf((size_t)al + 1);
      CHECK(cibh5r_read(r, ao, al, buf.data()) == al);
      if (std::string(an) == "model_config") CHECK(std::string(buf.data(), (size_t)al) == cfg);
      if (std::string(an) == "lr") { double v; memcpy(&v, buf.data(), 8); CHECK(v == lr); }
      if (std::string(an) == "layer_names") CHECK(ar == 1 && ash[0] == 3 && ai == 8 && memcmp(buf.data(), names, 24) == 0);
    }
    if (isg) { ++groups; continue; }
    ++datasets;
    std::vector<char> buf((size_t)len + 1);
    CHECK(cibh5r_read(r, off, len, buf.data()) == len);
    const std::string sp(p);
    if (sp.rfind("/model_weights/w", 0) == 0) {
      const int idx = atoi(sp.c_str() + 16);
      CHECK(idx >= 0 && idx < 150 && kind == 0 && isz == 4 && rank == 1 && shape[0] == idx + 1);
      CHECK(memcmp(buf.data(), payloads[idx].data(), (size_t)len) == 0);
    } else if (sp == "/optimizer_weights/iterations") {
      CHECK(kind == 1 && isz == 8 && rank == 2 && shape[0] == 3 && shape[1] == 5 && memcmp(buf.data(), ints.data(), 120) == 0);
    }
  }
  CHECK(datasets == 151 && groups == 4 && attrs == 3);
  cibh5r_close(r);
  // a file that is not HDF5 must be refused cleanly
  FILE* f = fopen((path + ".bad").c_str(), "wb"); CHECK(f); fputs("definitely not hdf5, but longer than a superblock ..............................................................", f); fclose(f);
  CHECK(cibh5r_open((path + ".bad").c_str()) == nullptr);
  remove((path + ".bad").c_str()); remove(path.c_str());
  printf("h5 round trip ok: %d datasets, %d groups, %d attributes\n", datasets, groups, attrs);
  return 0;
}
