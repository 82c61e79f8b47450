// Gradient all-reduce over NVLink 5 / NVSwitch peer memory, FUSED with the 1/world scale and the optimizer update.
//
// Every rank (one process or one device per GPU) owns an "arena" -- params | grads | slot0 | slot1 | flags -- that
// all peers have mapped (CUDA IPC or same-process peer access).  The kernel is launched with the same grid on every
// rank after that rank's wgrad kernels (stream order makes the local gradients final).
//
//   two-shot (default):  barrier -> each rank sums ITS 1/world slice of the gradient from all peers in fixed rank
//                        order -> x 1/world -> optimizer update of that slice (moments are thereby sharded, ZeRO-1
//                        style) -> writes the new parameters into every peer's parameter buffer (all-gather by peer
//                        stores) -> barrier -> clears its own gradient buffer.
//   one-shot (small n):  barrier -> every rank sums the whole gradient from all peers in fixed rank order and
//                        updates its own replica redundantly -> barrier -> clears its own gradient buffer.
//
// Fixed summation order => replicas stay bit-identical (reference invariant: DistTrain_mnist.ipynb:526-548 prints
// the same test loss on all 8 ranks).  Barriers are per-CTA flag words written with st.release.sys and polled with
// ld.acquire.sys; the flag value is a monotonically increasing epoch kept in device memory, so a captured CUDA graph
// can be replayed every step.  A bounded spin turns a dead peer into an error instead of a GPU hang.
//
// Replaces hvd.DistributedOptimizer (Horovod fusion buffer + MPI_Allreduce, reference rpv.py:63-65) and the
// optimizer's apply_gradients in one launch; no NCCL on this path.
#pragma once
#include "ptx.cuh"
#include "kernels_basic.cuh"

namespace cib {

#define CIB_MAX_RANKS 8
#define CIB_AR_MAX_CTAS 296

struct CommDev {                       // lives in kernel parameter space
  float* grads[CIB_MAX_RANKS];         // peer-mapped gradient buffers (index = rank)
  float* params[CIB_MAX_RANKS];
  unsigned int* flags[CIB_MAX_RANKS];  // [2 phases][CIB_AR_MAX_CTAS][CIB_MAX_RANKS]
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// All CTAs with the same blockIdx on all ranks meet here.  `phase` selects the flag bank.
__device__ __forceinline__ void peer_barrier(const CommDev& c, int phase, unsigned int epoch) {
  __syncthreads();
  if (threadIdx.x < c.world) {
    const int q = threadIdx.x;
    __threadfence_system();
    st_release_sys(c.flags[q] + ((size_t)phase * CIB_AR_MAX_CTAS + blockIdx.x) * CIB_MAX_RANKS + c.rank, epoch);
    const unsigned int* mine = c.flags[c.rank] + ((size_t)phase * CIB_AR_MAX_CTAS + blockIdx.x) * CIB_MAX_RANKS + q;
    long long t0 = clock64();
    // epochs only grow; "at least epoch" tolerates a peer that is already one step ahead
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) {
      if (clock64() - t0 > CIB_WAIT_TIMEOUT_CYCLES) dev_fail(0xAB000000u | (phase << 8) | q, epoch, *mine);
    }
  }
  __syncthreads();
}

struct AllreduceOptParams {
  CommDev comm;
  float* slot0; float* slot1;          // local optimizer state
  int n4;                              // float4 count of the flat buffers (multiple of world)
  int kind;                            // optimizer kind (see optimizer_kernel)
  int one_shot;
  HyperBlock* hb;
  unsigned int* epoch_ctr;             // device counter, advanced by the last block
  int advance;                         // 1: the step's last bucket (moves the optimizer's step counter on)
};

__global__ void __launch_bounds__(512) allreduce_opt_kernel(const AllreduceOptParams p) {
  const CommDev& c = p.comm;
  const int W = c.world, R = c.rank;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  const OptCoef coef = opt_coefficients(p.hb, p.kind);
  const float inv_w = 1.0f / (float)W;

  peer_barrier(c, 0, epoch);           // every rank's gradients are final

  // CTA b owns float4 range [lo, hi); in two-shot mode rank r handles the r-th 1/W of it
  const int per_cta = (p.n4 + gridDim.x - 1) / gridDim.x;
  const int lo = blockIdx.x * per_cta;
  const int hi = min(p.n4, lo + per_cta);
  int my_lo = lo, my_hi = hi;
  if (!p.one_shot) {
    const int span = (hi - lo + W - 1) / W;
    my_lo = min(hi, lo + R * span); my_hi = min(hi, my_lo + span);
  }
  float4* lp = reinterpret_cast<float4*>(c.params[R]);
  float4* s0 = reinterpret_cast<float4*>(p.slot0);
  float4* s1 = p.slot1 ? reinterpret_cast<float4*>(p.slot1) : nullptr;
  for (int i = my_lo + threadIdx.x; i < my_hi; i += blockDim.x) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int q0 = 0; q0 < W; q0 += 4) {          // up to 4 peer loads in flight, summed in rank order
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u < W) t[u] = ld_peer_f4(c.grads[q0 + u] + 4 * (size_t)i);
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u < W) { g.x += t[u].x; g.y += t[u].y; g.z += t[u].z; g.w += t[u].w; }
    }
    g.x *= inv_w; g.y *= inv_w; g.z *= inv_w; g.w *= inv_w;
    float4 pv = lp[i];
    float4 a = s0[i];
    float4 b = s1 ? s1[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    opt_update(p.kind, p.hb, coef, g.x, pv.x, a.x, b.x);
    opt_update(p.kind, p.hb, coef, g.y, pv.y, a.y, b.y);
    opt_update(p.kind, p.hb, coef, g.z, pv.z, a.z, b.z);
    opt_update(p.kind, p.hb, coef, g.w, pv.w, a.w, b.w);
    s0[i] = a;
    if (s1) s1[i] = b;
    if (p.one_shot) {
      lp[i] = pv;
    } else {
#pragma unroll 1
      for (int q = 0; q < W; ++q) st_peer_f4(c.params[(R + q) % W] + 4 * (size_t)i, pv);   // stagger targets
    }
  }

  peer_barrier(c, 1, epoch);           // all peers have read my gradients / written my parameters

  float4* lg = reinterpret_cast<float4*>(c.grads[R]);
  for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) lg[i] = make_float4(0.f, 0.f, 0.f, 0.f);

  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&p.hb->ticket, 1u);
    if (tk == gridDim.x - 1) { p.hb->ticket = 0; *p.epoch_ctr = epoch; if (p.advance) opt_advance_state(p.hb, p.kind); __threadfence(); }
  }
}

// Plain fused all-reduce (sum or average) of a small/large buffer, no optimizer: used for the bandwidth sweep,
// MetricAverage (reference rpv.py:87) and for cross-checking against NCCL.  Result lands in every rank's buffer.
struct AllreduceParams {
  CommDev comm;            // .grads = the buffers to reduce in place; .params unused
  int n4; int average; int one_shot;
  unsigned int* epoch_ctr; unsigned int* ticket;
  float* scratch;          // one-shot: local staging [n4*4] so in-place results do not race with peers' reads
};

__global__ void __launch_bounds__(512) allreduce_kernel(const AllreduceParams p) {
  const CommDev& c = p.comm;
  const int W = c.world, R = c.rank;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  const float sc = p.average ? 1.0f / (float)W : 1.0f;
  peer_barrier(c, 0, epoch);
  const int per_cta = (p.n4 + gridDim.x - 1) / gridDim.x;
  const int lo = blockIdx.x * per_cta, hi = min(p.n4, lo + per_cta);
  int my_lo = lo, my_hi = hi;
  if (!p.one_shot) { const int span = (hi - lo + W - 1) / W; my_lo = min(hi, lo + R * span); my_hi = min(hi, my_lo + span); }
  for (int i = my_lo + threadIdx.x; i < my_hi; i += blockDim.x) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int q0 = 0; q0 < W; q0 += 4) {
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u < W) t[u] = ld_peer_f4(c.grads[q0 + u] + 4 * (size_t)i);
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u < W) { g.x += t[u].x; g.y += t[u].y; g.z += t[u].z; g.w += t[u].w; }
    }
    g.x *= sc; g.y *= sc; g.z *= sc; g.w *= sc;
    if (p.one_shot) reinterpret_cast<float4*>(p.scratch)[i] = g;
    else {
      // slices are disjoint across ranks, so writing results straight into every peer's buffer is race-free
#pragma unroll 1
      for (int q = 0; q < W; ++q) st_peer_f4(c.grads[(R + q) % W] + 4 * (size_t)i, g);
    }
  }
  peer_barrier(c, 1, epoch);
  if (p.one_shot) {
    float4* lg = reinterpret_cast<float4*>(c.grads[R]);
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) lg[i] = reinterpret_cast<const float4*>(p.scratch)[i];
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(p.ticket, 1u);
    if (tk == gridDim.x - 1) { *p.ticket = 0; *p.epoch_ctr = epoch; __threadfence(); }
  }
}

// Broadcast from `root` (reference rpv.py:85 BroadcastGlobalVariablesCallback(0)): every rank pulls root's buffer.
__global__ void __launch_bounds__(512) broadcast_kernel(const AllreduceParams p, int root) {
  const CommDev& c = p.comm;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  peer_barrier(c, 0, epoch);
  if (c.rank != root) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.n4; i += gridDim.x * blockDim.x)
      reinterpret_cast<float4*>(c.grads[c.rank])[i] = ld_peer_f4(c.grads[root] + 4 * (size_t)i);
  }
  peer_barrier(c, 1, epoch);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(p.ticket, 1u);
    if (tk == gridDim.x - 1) { *p.ticket = 0; *p.epoch_ctr = epoch; __threadfence(); }
  }
}

}  // namespace cib
