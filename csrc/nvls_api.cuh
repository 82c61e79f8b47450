// C ABI of the NVLS (NVSwitch multicast) all-reduce, see nvls.cuh (collective C1 of SURVEY 2.7.2: hvd.DistributedOptimizer, rpv.py:65).
#pragma once
#include "nvls.cuh"

using namespace cib;

extern "C" {

int cib_nvls_supported(int device, int* supported) {
  if (g_device < 0) CIB_TRY(cib_init(device));
  return nvls_supported(device, supported);
}
int cib_nvls_create_local(cib_nvls_t* out, int world, long long nfloats) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
  int cur = 0; cudaGetDevice(&cur);
  Nvls* s = nullptr;
  const int rc = nvls_create_local(&s, world, nfloats);
  cudaSetDevice(cur);
  if (rc) return rc;
  *out = s;
  return 0;
}
int cib_nvls_ptr(cib_nvls_t s, int rank, void** dev_ptr) { *dev_ptr = (void*)reinterpret_cast<Nvls*>(s)->uc[rank]; return 0; }
int cib_nvls_allreduce(cib_nvls_t s, int rank, long long nfloats, int average, void* stream) {
  return nvls_allreduce(reinterpret_cast<Nvls*>(s), rank, nfloats, average, (cudaStream_t)stream);
}
int cib_nvls_destroy(cib_nvls_t s) {
  std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
  return nvls_destroy(reinterpret_cast<Nvls*>(s));
}

}  // extern "C"
