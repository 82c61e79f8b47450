// Thin inline-PTX layer for sm_100a: mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (MMA/TMEM).
// (The reference has no native code at all -- SURVEY 2.7: "zero .cu files, zero PTX" -- so there is nothing this mirrors.)
// Everything here is hand-written for Blackwell; nothing is shared with any other architecture.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace cib {

// ---------------------------------------------------------------------------------------------
// Error flag written by device code before it traps (bounded waits instead of hangs).
// ---------------------------------------------------------------------------------------------
__device__ unsigned int g_dev_error[4];   // [0]=code, [1]=blockIdx, [2]=aux0, [3]=aux1

#ifndef CIB_SPIN_BEFORE_SLEEP
#define CIB_SPIN_BEFORE_SLEEP 4096
#endif
#ifndef CIB_WAIT_TIMEOUT_CYCLES
#define CIB_WAIT_TIMEOUT_CYCLES (4000000000LL)   // ~2 s at 1.9 GHz: a lost arrive becomes an error, not a hang
#endif

__device__ __forceinline__ void dev_fail(unsigned code, unsigned a0 = 0, unsigned a1 = 0) {
  g_dev_error[0] = code; g_dev_error[1] = blockIdx.x; g_dev_error[2] = a0; g_dev_error[3] = a1;
  __threadfence_system();
  __trap();
}

// Soft variant for waits on ANOTHER GPU (a dead or late peer process is an operational condition, not a kernel bug): the first
// failure is recorded, the waiting thread gives up and the kernel runs to its end on whatever data is there; the host sees the code
// at its next cib_engine_sync / cib_engine_get_metrics and fails the call.  No trap, so the CUDA context stays usable.
__device__ __forceinline__ void dev_fail_soft(unsigned code, unsigned a0 = 0, unsigned a1 = 0) {
  if (atomicCAS(&g_dev_error[0], 0u, code) == 0u) { g_dev_error[1] = blockIdx.x; g_dev_error[2] = a0; g_dev_error[3] = a1; }
  __threadfence_system();
}

// Programmatic dependent launch (griddepcontrol): a kernel launched with the programmatic-stream-serialization attribute may start
// while its predecessor in the stream is still running.  pdl_launch() at the top of EVERY kernel lets the successor's CTAs be
// scheduled as soon as all of this grid's CTAs have started; pdl_wait() blocks until the predecessor grid has completed and its
// memory is visible -- it is placed after the on-chip prologue (mbarrier init, TMEM allocation, descriptor prefetch, index tables)
// and before the first access to global memory that another kernel of the step produces or consumes.  Both are no-ops in a
// kernel launched the ordinary way.
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Optional in-kernel timeline: when g_ktrace points at a buffer, role leaders of CTA 0 and CTA 1 drop globaltimer (ns) stamps.
__device__ unsigned long long* g_ktrace = nullptr;
__device__ __forceinline__ void ktrace(int slot) {
  unsigned long long* t = g_ktrace;
  if (t != nullptr && blockIdx.x < 2) {
    unsigned long long ns;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
    t[blockIdx.x * 32 + slot] = ns;
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "elect.sync _|P1, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done != 0;
}
// Bounded wait: traps (with a code in g_dev_error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, unsigned who) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  int spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    // Back off only on LONG waits: a sleep in the steady state of a producer/consumer pipeline adds its wake-up latency to
    // every tile (measured: 0.75 us per tile of pure hand-shake with a 32 ns sleep after 4 tries).
    if (++spins > CIB_SPIN_BEFORE_SLEEP) __nanosleep(64);
    if ((spins & 1023) == 0 && clock64() - t0 > CIB_WAIT_TIMEOUT_CYCLES) dev_fail(0xBA000000u | who, bar, parity);
  }
}

// ---------------------------------------------------------------------------------------------
// TMA: tiled tensor-map loads, completion signalled on an mbarrier (complete_tx::bytes).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tm, uint32_t bar,
                                            int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05: TMEM allocation, MMA, commit, TMEM loads
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {   // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {    // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], tf32 inputs, fp32 accumulate. One thread issues.
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// All previously issued MMAs of this thread arrive on the mbarrier when they complete
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {           // 32 lanes x 16 columns
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// Descriptors.
// Shared-memory matrix descriptor (64 bit): start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) |
// version=1 [46,48) | base_offset [49,52) | layout [61,64)  (2 = SWIZZLE_128B, 4 = SWIZZLE_64B)
// ---------------------------------------------------------------------------------------------
// SWZ_128_B32 = 128-byte swizzle with a 32-byte base (Swizzle<2,5,2>, 4-row atoms): the ONLY layout the tensor core
// accepts for MN-major tf32 operands; it pairs with the TMA mode CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
enum : uint32_t { SWZ_128 = 2, SWZ_64 = 4, SWZ_32 = 6, SWZ_NONE = 0, SWZ_128_B32 = 1 };

__device__ __forceinline__ uint64_t make_sdesc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout & 7) << 61;
  return d;
}
// Instruction descriptor for kind::tf32 with fp32 accumulation.
// c_format=F32 [4,6) | a_format=TF32 [7,10) | b_format=TF32 [10,13) | a_major [15] | b_major [16] | N>>3 [17,23) | M>>4 [24,29)
__host__ __device__ __forceinline__ uint32_t make_idesc_tf32(uint32_t M, uint32_t N, uint32_t a_mn_major, uint32_t b_mn_major) {
  uint32_t d = 0;
  d |= 1u << 4;
  d |= 2u << 7;
  d |= 2u << 10;
  d |= (a_mn_major & 1u) << 15;
  d |= (b_mn_major & 1u) << 16;
  d |= ((N >> 3) & 0x3Fu) << 17;
  d |= ((M >> 4) & 0x1Fu) << 24;
  return d;
}

// ---------------------------------------------------------------------------------------------
// Small helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void red_add_f32(float* addr, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add_v4_f32(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Counter-based RNG for dropout (graph-replay safe: keyed by seed, step, layer, element).
__device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ float dropout_uniform(uint32_t seed, uint32_t step, uint32_t layer, uint32_t idx) {
  uint32_t h = hash_u32(idx ^ hash_u32(seed + 0x9E3779B9u * (step + 1u)) ^ (layer * 0x85EBCA6Bu));
  h = hash_u32(h + layer + 0x27D4EB2Fu);
  return (float)(h >> 8) * (1.0f / 16777216.0f);   // [0,1)
}

}  // namespace cib
