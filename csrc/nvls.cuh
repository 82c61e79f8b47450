// NVLS all-reduce: the reduction happens INSIDE the NVSwitch (multimem.ld_reduce) and the result is broadcast by the switch
// (multimem.st), so every GPU moves 1/N of the buffer each way instead of pulling (N-1)/N of it from its peers.
//
// Set-up (CUDA virtual memory management + multicast objects, driver API): one physical allocation per GPU, bound at the
// same offset of one multicast object; every GPU maps (a) all the unicast allocations (peer flags for the barrier) and
// (b) the multicast object.  This version serves communicators whose GPUs live in ONE process (cib_comm_create_local's
// case: several engines of one process, the single-process multi-GPU checks); across processes the handles would have to
// travel as POSIX file descriptors, which is left for a later round.  Where the platform does not expose multicast
// (attribute CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 0, or the fabric manager has NVLS disabled) creation fails with a
// clear message and the peer-to-peer kernels of allreduce.cuh are used.
//
// Collective call site in the reference: hvd.DistributedOptimizer (rpv.py:65) -- SURVEY 2.7.2 C1, "NVLS variant".
#pragma once
#include "engine_graph.cuh"

namespace cib {

struct Nvls {
  int world = 0;
  size_t data_bytes = 0, bytes = 0;               // payload / whole allocation (payload + flags + counters)
  size_t off_flags = 0, off_ctr = 0;
  CUmemGenericAllocationHandle mc = 0;
  CUmemGenericAllocationHandle mem[CIB_MAX_RANKS] = {};
  CUdeviceptr uc[CIB_MAX_RANKS] = {};             // unicast mapping of rank r's memory (readable from every GPU)
  CUdeviceptr mcva = 0;                           // one multicast mapping, accessible from every GPU
  bool bound[CIB_MAX_RANKS] = {};
};

struct NvlsParams {
  CommDev comm;                 // .flags[q] = barrier flags of rank q (unicast); .grads / .params unused
  float* mc;                    // multicast pointer
  int n4; int average;
  unsigned int* epoch_ctr; unsigned int* ticket;
};

__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Rank r reduces the r-th 1/W of every CTA's range in the switch and broadcasts it to all replicas through the switch.
__global__ void __launch_bounds__(512) nvls_allreduce_kernel(const NvlsParams p) {
  const CommDev& c = p.comm;
  const int W = c.world, R = c.rank;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  const float sc = p.average ? 1.0f / (float)W : 1.0f;
  peer_barrier(c, 0, epoch);                                    // every rank's input is in place
  const int per_cta = (p.n4 + gridDim.x - 1) / gridDim.x;
  const int lo = blockIdx.x * per_cta, hi = min(p.n4, lo + per_cta);
  const int span = (hi - lo + W - 1) / W;
  const int my_lo = min(hi, lo + R * span), my_hi = min(hi, my_lo + span);
  for (int i = my_lo + threadIdx.x; i < my_hi; i += blockDim.x) {
    float4 v = multimem_ld_reduce_f4(p.mc + 4 * (size_t)i);
    v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;
    multimem_st_f4(p.mc + 4 * (size_t)i, v);
  }
  __threadfence_system();
  peer_barrier(c, 1, epoch);                                    // all slices have been written everywhere
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(p.ticket, 1u);
    if (tk == gridDim.x - 1) { *p.ticket = 0; *p.epoch_ctr = epoch; __threadfence(); }
  }
}

static int nvls_supported(int device, int* out) {
  *out = 0;
  CIB_TRY(nvls_load_driver());
  CUdevice d; CIB_CU(g_nvls.DeviceGet(&d, device));
  int v = 0;
  CIB_CU(g_nvls.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d));
  *out = v;
  return 0;
}

static int nvls_destroy(Nvls* s) {
  if (!s) return 0;
  if (g_nvls.ok) {
    if (s->mcva) { g_nvls.MemUnmap(s->mcva, s->bytes); g_nvls.MemAddressFree(s->mcva, s->bytes); }
    for (int r = 0; r < s->world; ++r) {
      if (s->bound[r]) { CUdevice d; if (g_nvls.DeviceGet(&d, r) == CUDA_SUCCESS) g_nvls.MulticastUnbind(s->mc, d, 0, s->bytes); }
      if (s->uc[r]) { g_nvls.MemUnmap(s->uc[r], s->bytes); g_nvls.MemAddressFree(s->uc[r], s->bytes); }
      if (s->mem[r]) g_nvls.MemRelease(s->mem[r]);
    }
    if (s->mc) g_nvls.MemRelease(s->mc);
  }
  delete s;
  return 0;
}

// GPUs 0 .. world-1 of this process.
static int nvls_create_local(Nvls** out, int world, long long nfloats) {
  if (world < 2 || world > CIB_MAX_RANKS) return fail("nvls: world must be 2..8");
  CIB_TRY(nvls_load_driver());
  for (int r = 0; r < world; ++r) {
    int sup = 0; CIB_TRY(nvls_supported(r, &sup));
    if (!sup) return fail("nvls: device " + std::to_string(r) + " reports CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 0");
  }
  for (int r = 0; r < world; ++r) { CIB_CUDA(cudaSetDevice(r)); CIB_CUDA(cudaFree(0)); }     // make sure every context exists
  Nvls* s = new Nvls(); s->world = world;
  s->data_bytes = align_up((size_t)nfloats * 4, 1024);
  s->off_flags = s->data_bytes;
  s->off_ctr = s->off_flags + align_up((size_t)2 * CIB_AR_MAX_CTAS * CIB_MAX_RANKS * 4, 1024);
  size_t need = s->off_ctr + 1024;

  CUmulticastObjectProp mp; memset(&mp, 0, sizeof mp);
  mp.numDevices = (unsigned)world; mp.handleTypes = 0; mp.flags = 0; mp.size = need;
  size_t gran = 0;
  CUresult r0 = g_nvls.MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM);
  if (r0 != CUDA_SUCCESS || gran == 0) { delete s; return fail("nvls: cuMulticastGetGranularity -> CUresult " + std::to_string((int)r0)); }
  s->bytes = align_up(need, gran);
  mp.size = s->bytes;
  CUresult rc = g_nvls.MulticastCreate(&s->mc, &mp);
  if (rc != CUDA_SUCCESS) { s->mc = 0; nvls_destroy(s); return fail("nvls: cuMulticastCreate -> CUresult " + std::to_string((int)rc) + " (multicast not enabled on this platform)"); }
  auto bail = [&](const std::string& m) { nvls_destroy(s); return fail(m); };
  for (int r = 0; r < world; ++r) {
    CUdevice d; if (g_nvls.DeviceGet(&d, r) != CUDA_SUCCESS) return bail("nvls: cuDeviceGet");
    CUresult e = g_nvls.MulticastAddDevice(s->mc, d);
    if (e != CUDA_SUCCESS) return bail("nvls: cuMulticastAddDevice -> CUresult " + std::to_string((int)e));
  }
  for (int r = 0; r < world; ++r) {
    CUmemAllocationProp ap; memset(&ap, 0, sizeof ap);
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED; ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ap.location.id = r;
    CUresult e = g_nvls.MemCreate(&s->mem[r], s->bytes, &ap, 0);
    if (e != CUDA_SUCCESS) { s->mem[r] = 0; return bail("nvls: cuMemCreate -> CUresult " + std::to_string((int)e)); }
    e = g_nvls.MulticastBindMem(s->mc, 0, s->mem[r], 0, s->bytes, 0);
    if (e != CUDA_SUCCESS) return bail("nvls: cuMulticastBindMem -> CUresult " + std::to_string((int)e));
    s->bound[r] = true;
  }
  std::vector<CUmemAccessDesc> acc(world);
  for (int r = 0; r < world; ++r) { acc[r].location.type = CU_MEM_LOCATION_TYPE_DEVICE; acc[r].location.id = r; acc[r].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE; }
  for (int r = 0; r < world; ++r) {
    CUresult e = g_nvls.MemAddressReserve(&s->uc[r], s->bytes, gran, 0, 0);
    if (e != CUDA_SUCCESS) { s->uc[r] = 0; return bail("nvls: cuMemAddressReserve -> CUresult " + std::to_string((int)e)); }
    e = g_nvls.MemMap(s->uc[r], s->bytes, 0, s->mem[r], 0);
    if (e != CUDA_SUCCESS) return bail("nvls: cuMemMap(unicast) -> CUresult " + std::to_string((int)e));
    e = g_nvls.MemSetAccess(s->uc[r], s->bytes, acc.data(), (size_t)world);
    if (e != CUDA_SUCCESS) return bail("nvls: cuMemSetAccess(unicast) -> CUresult " + std::to_string((int)e));
  }
  {
    CUresult e = g_nvls.MemAddressReserve(&s->mcva, s->bytes, gran, 0, 0);
    if (e != CUDA_SUCCESS) { s->mcva = 0; return bail("nvls: cuMemAddressReserve(mc) -> CUresult " + std::to_string((int)e)); }
    e = g_nvls.MemMap(s->mcva, s->bytes, 0, s->mc, 0);
    if (e != CUDA_SUCCESS) return bail("nvls: cuMemMap(multicast) -> CUresult " + std::to_string((int)e));
    e = g_nvls.MemSetAccess(s->mcva, s->bytes, acc.data(), (size_t)world);
    if (e != CUDA_SUCCESS) return bail("nvls: cuMemSetAccess(multicast) -> CUresult " + std::to_string((int)e));
  }
  int cur = 0; cudaGetDevice(&cur);
  for (int r = 0; r < world; ++r) {
    if (cudaSetDevice(r) != cudaSuccess || cudaMemset((void*)s->uc[r], 0, s->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
      cudaSetDevice(cur); return bail("nvls: clearing the buffers failed");
    }
  }
  cudaSetDevice(cur);
  *out = s;
  return 0;
}

static int nvls_allreduce(Nvls* s, int rank, long long nfloats, int average, cudaStream_t st) {
  if (nfloats % 4 || (size_t)nfloats * 4 > s->data_bytes) return fail("nvls: size must be a multiple of 4 floats within the buffer");
  NvlsParams p; memset(&p, 0, sizeof p);
  p.comm.rank = rank; p.comm.world = s->world; p.comm.timeout_ns = g_peer_timeout_ns;
  for (int q = 0; q < s->world; ++q) p.comm.flags[q] = (unsigned int*)(s->uc[q] + s->off_flags);
  p.mc = (float*)s->mcva;
  p.n4 = (int)(nfloats / 4); p.average = average;
  p.epoch_ctr = (unsigned int*)(s->uc[rank] + s->off_ctr); p.ticket = p.epoch_ctr + 1;
  CIB_CUDA(cudaSetDevice(rank));
  nvls_allreduce_kernel<<<ar_grid_for(p.n4), 512, 0, st>>>(p);
  CIB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace cib
