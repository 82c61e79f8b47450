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

// Time-out of every cross-GPU wait, in nanoseconds of %globaltimer.  Ranks are separate processes: the host-side skew between them
// (dataset upload, graph capture, a rank-0-only checkpoint write) is seconds, not the microseconds an intra-kernel mbarrier wait
// sees, so this is minutes by default (CIB_PEER_TIMEOUT_S overrides it; cib_set_peer_timeout_s at run time).  A dead peer still
// becomes an error code in cib_device_error (and a failed cib_engine_sync) instead of a hang.
static unsigned long long g_peer_timeout_ns = 180ull * 1000000000ull;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long ns;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
  return ns;
}

struct CommDev {                       // lives in kernel parameter space
  float* grads[CIB_MAX_RANKS];         // peer-mapped gradient buffers (index = rank)
  float* params[CIB_MAX_RANKS];
  unsigned int* flags[CIB_MAX_RANKS];  // [2 phases][CIB_AR_MAX_CTAS][CIB_MAX_RANKS]
  int rank, world;
  unsigned long long timeout_ns;       // see g_peer_timeout_ns
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
    const unsigned long long t0 = globaltimer_ns();
    unsigned int spins = 0;
    // epochs only grow; "at least epoch" tolerates a peer that is already one step ahead
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) {
      if ((++spins & 1023u) == 0u && globaltimer_ns() - t0 > c.timeout_ns) { dev_fail_soft(0xAB000000u | (phase << 8) | q, epoch, *mine); break; }
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

// =================================================================================================================
// PUSH variant of the fused gradient all-reduce + optimizer (the default on the training path).
//
// The pull kernel above needs two full barriers per launch and keeps hundreds of threads stalled on NVLink round trips; that
// made it both slow for small buckets (two barriers = ~15 us of pure latency) and a bad neighbour when it runs next to the
// convolution backward.  Here every rank STORES its gradients into a receive area of the consumer (stores are fire and
// forget: few threads move the whole bucket), raises one flag per CTA and peer, and then works on LOCAL memory only:
//
//   two-shot (large bucket): rank r owns the r-th 1/W slice.  Every rank pushes slice q of its gradients into rank q's receive
//                        slot [sender]; flag; wait for the W-1 flags; sum the W contributions of the own slice in rank order
//                        (the local one straight from the gradient buffer); x 1/W; optimizer on the slice (moments sharded,
//                        cib_engine_gather_slots reassembles them); push the new parameters to every peer; flag; wait.
//   one-shot (small bucket): every rank pushes the whole bucket to every peer; flag; wait; every rank sums all W contributions in
//                        rank order and updates its full replica.  ONE one-way flag round, no second barrier at all.
//
// No barrier is needed before a push: a receive slot is only rewritten by the same peer's next collective of the same bucket,
// and that peer cannot get there before this rank has finished reading (its next collective waits for a flag that this rank
// raises only from a later kernel of the same stream).  The sender clears its own gradients itself right after pushing --
// nobody reads them remotely -- so there is no "all peers are done reading" barrier either.
// CTA j of every rank handles sub-range j of every slice, and the flags are per (CTA, sender): no grid-wide dependency, so
// CTAs need not be co-resident.  128 threads and <= 72 registers per CTA: a CTA fits on an SM NEXT to a 576-thread x 96-register
// convolution CTA, so the early bucket really overlaps with the convolution backward instead of displacing it.
// Fixed summation order => replicas stay bit-identical.
// Replaces hvd.DistributedOptimizer (reference rpv.py:63-65) + apply_gradients in one launch per bucket.
// =================================================================================================================
#define CIB_PAR_MAX_CTAS 148
#define CIB_PAR_THREADS 128
#define CIB_PAR_UNROLL 4
#define CIB_LL_MAX4 16384               // largest bucket (float4) that takes the LL protocol: 256 KB

struct PushArParams {
  float* grads;                        // local gradient bucket
  float* params_l;                     // local parameter bucket
  float* slot0; float* slot1;          // local optimizer state of the bucket
  float* recv[CIB_MAX_RANKS];          // receive area of rank q for this bucket (recv[rank] is local): [W senders][S4] float4;
                                       // one-shot: two such areas, used alternately (epoch parity) -- a peer that is one
                                       // collective ahead must not overwrite what this rank is still summing
  float* params[CIB_MAX_RANKS];        // parameter bucket of rank q (two-shot all-gather target)
  unsigned int* flags[CIB_MAX_RANKS];  // flag table of rank q for this bucket: [2 phases][CIB_PAR_MAX_CTAS][CIB_MAX_RANKS]
  int rank, world;
  int n4;                              // float4 count of the bucket
  int S4;                              // two-shot: slice length ceil(n4 / world); one-shot: n4
  int one_shot;
  int kind, advance;
  HyperBlock* hb;
  unsigned int* epoch_ctr;             // this bucket's collective counter (device memory, advanced by the last CTA)
  unsigned int* ticket;
  unsigned long long timeout_ns;
  unsigned long long* trace;           // optional: [CTA 0 only] globaltimer stamps: start, flags seen, reduced, params seen
};

__device__ __forceinline__ void par_signal_and_wait(const PushArParams& p, int phase, unsigned int epoch) {
  __syncthreads();                     // every thread's stores of this CTA are ordered before the release below
  if (threadIdx.x < p.world && (int)threadIdx.x != p.rank) {
    const int q = threadIdx.x;
    const size_t slot = ((size_t)phase * CIB_PAR_MAX_CTAS + blockIdx.x) * CIB_MAX_RANKS;
    __threadfence_system();
    st_release_sys(p.flags[q] + slot + p.rank, epoch);
    const unsigned int* mine = p.flags[p.rank] + slot + q;
    const unsigned long long t0 = globaltimer_ns();
    unsigned int spins = 0;
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) {
      if ((++spins & 1023u) == 0u && globaltimer_ns() - t0 > p.timeout_ns) { dev_fail_soft(0xAC000000u | (phase << 8) | q, epoch, *mine); break; }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_cg_f4(const float4* p) { return __ldcg(p); }

__global__ void __launch_bounds__(CIB_PAR_THREADS, 8) push_allreduce_opt_kernel(const PushArParams p) {
  const int W = p.world, R = p.rank, G = gridDim.x, j = blockIdx.x;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  const bool tr = p.trace != nullptr && j == 0 && threadIdx.x == 0;
  if (tr) p.trace[0] = globaltimer_ns();
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4* lg = reinterpret_cast<float4*>(p.grads);
  const size_t par_off = p.one_shot ? (size_t)(epoch & 1u) * (size_t)W * p.S4 : 0;
  constexpr int U = CIB_PAR_UNROLL, T = CIB_PAR_THREADS;

  // ---- 1. push: sub-range j of slice q goes to rank q (one-shot: the one "slice" is the whole bucket, for every q).
  //      U independent loads per thread are in flight before the first store (the loop is latency-, not bandwidth-bound).
  if (p.one_shot) {
    const int per = (p.n4 + G - 1) / G;
    const int lo = min(p.n4, j * per), hi = min(p.n4, lo + per);
    for (int i0 = lo + threadIdx.x; i0 < hi; i0 += U * T) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) if (i0 + u * T < hi) v[u] = lg[i0 + u * T];
      for (int d = 1; d < W; ++d) {
        float* dst = p.recv[(R + d) % W] + 4 * ((size_t)R * p.S4 + par_off);
#pragma unroll
        for (int u = 0; u < U; ++u) if (i0 + u * T < hi) st_peer_f4(dst + 4 * (size_t)(i0 + u * T), v[u]);
      }
    }
  } else {
    for (int d = 1; d < W; ++d) {
      const int q = (R + d) % W;                                 // stagger the targets over the ranks
      const int base = q * p.S4;
      const int len = max(0, min(p.n4, base + p.S4) - base);
      const int per = (len + G - 1) / G;
      const int lo = min(len, j * per), hi = min(len, lo + per);
      float* dst = p.recv[q] + 4 * ((size_t)R * p.S4);
      for (int i0 = lo + threadIdx.x; i0 < hi; i0 += U * T) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) if (i0 + u * T < hi) v[u] = lg[base + i0 + u * T];
#pragma unroll
        for (int u = 0; u < U; ++u) if (i0 + u * T < hi) {
          st_peer_f4(dst + 4 * (size_t)(i0 + u * T), v[u]);
          lg[base + i0 + u * T] = zero4;                         // nobody else reads my gradients: cleared as soon as they are sent
        }
      }
    }
  }
  par_signal_and_wait(p, 0, epoch);
  if (tr) p.trace[1] = globaltimer_ns();

  // ---- 2. reduce my slice in rank order from LOCAL memory, x 1/W, optimizer
  const OptCoef coef = opt_coefficients(p.hb, p.kind);
  const float inv_w = 1.0f / (float)W;
  const int base = p.one_shot ? 0 : R * p.S4;
  const int len = max(0, min(p.n4, base + p.S4) - base);
  const int per = (len + G - 1) / G;
  const int lo = min(len, j * per), hi = min(len, lo + per);
  float4* lp = reinterpret_cast<float4*>(p.params_l);
  float4* s0 = reinterpret_cast<float4*>(p.slot0);
  float4* s1 = p.slot1 ? reinterpret_cast<float4*>(p.slot1) : nullptr;
  const float4* rbase = reinterpret_cast<const float4*>(p.recv[R]) + par_off;
  for (int i = lo + threadIdx.x; i < hi; i += T) {
    // loads first: parameter and moments, then the W contributions four at a time (the local one from the gradient buffer)
    float4 pv = lp[base + i];
    float4 a = s0[base + i];
    float4 b = s1 ? s1[base + i] : zero4;
    float4 g = zero4;
#pragma unroll 1
    for (int q0 = 0; q0 < W; q0 += 4) {
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (q0 + u < W) t[u] = (q0 + u) == R ? lg[base + i] : ld_cg_f4(rbase + (size_t)(q0 + u) * p.S4 + i);
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u < W) { g.x += t[u].x; g.y += t[u].y; g.z += t[u].z; g.w += t[u].w; }
    }
    lg[base + i] = zero4;
    g.x *= inv_w; g.y *= inv_w; g.z *= inv_w; g.w *= inv_w;
    opt_update(p.kind, p.hb, coef, g.x, pv.x, a.x, b.x);
    opt_update(p.kind, p.hb, coef, g.y, pv.y, a.y, b.y);
    opt_update(p.kind, p.hb, coef, g.z, pv.z, a.z, b.z);
    opt_update(p.kind, p.hb, coef, g.w, pv.w, a.w, b.w);
    s0[base + i] = a;
    if (s1) s1[base + i] = b;
    lp[base + i] = pv;
    if (!p.one_shot) {
#pragma unroll 1
      for (int d = 1; d < W; ++d) st_peer_f4(p.params[(R + d) % W] + 4 * (size_t)(base + i), pv);
    }
  }
  if (tr) p.trace[2] = globaltimer_ns();

  // ---- 3. two-shot: every peer's slice of the new parameters has landed here
  if (!p.one_shot) par_signal_and_wait(p, 1, epoch);
  if (tr) p.trace[3] = globaltimer_ns();

  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(p.ticket, 1u);
    if (tk == (unsigned int)G - 1u) { *p.ticket = 0u; *p.epoch_ctr = epoch; if (p.advance) opt_advance_state(p.hb, p.kind); __threadfence(); }
  }
}

// Latency-optimal one-shot variant for SMALL buckets ("LL": every 8 bytes that cross NVLink carry 4 bytes of data and a 4-byte
// flag, so the data are their own arrival signal).  The flag-based kernel above needs: stores -> wait for their acknowledgements
// (fence, one NVLink round trip) -> flag store -> poll; here a rank just stores {d0, epoch, d1, epoch} packets and the receiver spins
// on the packets themselves: ONE one-way trip, no fence, no flag table, no __syncthreads.  Costs twice the bytes, which is nothing for
// the ~100 KB convolution bucket this is used for (the last, un-hidden bucket of the step).  Receive area per rank:
// [2 parities][W senders][n4] x 32 bytes; the parity (epoch & 1) double-buffers against a peer that is one collective ahead, and the
// flag value (the epoch itself) tells a fresh packet from the one two collectives ago.
__device__ __forceinline__ void st_ll(float* dst16, float a, float b, unsigned int flag) {
  asm volatile("st.volatile.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst16), "r"(__float_as_uint(a)), "r"(flag), "r"(__float_as_uint(b)), "r"(flag) : "memory");
}
__device__ __forceinline__ bool ld_ll(const float* src16, float& a, float& b, unsigned int flag) {
  unsigned int x, f0, y, f1;
  asm volatile("ld.volatile.global.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(x), "=r"(f0), "=r"(y), "=r"(f1) : "l"(src16) : "memory");
  a = __uint_as_float(x); b = __uint_as_float(y);
  return f0 == flag && f1 == flag;
}

__global__ void __launch_bounds__(CIB_PAR_THREADS) ll_allreduce_opt_kernel(const PushArParams p) {
  const int W = p.world, R = p.rank, G = gridDim.x, j = blockIdx.x;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  const bool tr = p.trace != nullptr && j == 0 && threadIdx.x == 0;
  if (tr) p.trace[0] = globaltimer_ns();
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4* lg = reinterpret_cast<float4*>(p.grads);
  float4* lp = reinterpret_cast<float4*>(p.params_l);
  float4* s0 = reinterpret_cast<float4*>(p.slot0);
  float4* s1 = p.slot1 ? reinterpret_cast<float4*>(p.slot1) : nullptr;
  const size_t par = (size_t)(epoch & 1u) * W;
  const OptCoef coef = opt_coefficients(p.hb, p.kind);
  const float inv_w = 1.0f / (float)W;
  const int per = (p.n4 + G - 1) / G;
  const int lo = min(p.n4, j * per), hi = min(p.n4, lo + per);
  for (int i = lo + threadIdx.x; i < hi; i += CIB_PAR_THREADS) {
    const float4 mine = lg[i];
    for (int d = 1; d < W; ++d) {
      float* dst = p.recv[(R + d) % W] + 8 * ((par + R) * (size_t)p.n4 + i);
      st_ll(dst, mine.x, mine.y, epoch);
      st_ll(dst + 4, mine.z, mine.w, epoch);
    }
    float4 pv = lp[i];                                  // local loads overlap with the packets' flight
    float4 a = s0[i];
    float4 b = s1 ? s1[i] : zero4;
    lg[i] = zero4;
    // Poll ALL peers' packets in one round (independent loads: one L2 round trip per round, not one per peer), repeat for the ones
    // that have not arrived yet; the sum is formed afterwards in rank order whatever the arrival order was.
    float4 t[CIB_MAX_RANKS];
    unsigned int pending = ((1u << W) - 1u) & ~(1u << R);
    const unsigned long long t0 = globaltimer_ns();
    unsigned int spins = 0;
    while (pending) {
#pragma unroll
      for (int q = 0; q < CIB_MAX_RANKS; ++q) {
        if (pending & (1u << q)) {
          const float* src = p.recv[R] + 8 * ((par + q) * (size_t)p.n4 + i);
          const bool ok = ld_ll(src, t[q].x, t[q].y, epoch) & ld_ll(src + 4, t[q].z, t[q].w, epoch);
          if (ok) pending &= ~(1u << q);
        }
      }
      if (pending && (++spins & 1023u) == 0u && globaltimer_ns() - t0 > p.timeout_ns) { dev_fail_soft(0xAD000000u | pending, epoch, i); break; }
    }
    float4 g = zero4;
#pragma unroll
    for (int q = 0; q < CIB_MAX_RANKS; ++q) {
      if (q < W) {
        const float4 v = q == R ? mine : t[q];
        g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
      }
    }
    g.x *= inv_w; g.y *= inv_w; g.z *= inv_w; g.w *= inv_w;
    opt_update(p.kind, p.hb, coef, g.x, pv.x, a.x, b.x);
    opt_update(p.kind, p.hb, coef, g.y, pv.y, a.y, b.y);
    opt_update(p.kind, p.hb, coef, g.z, pv.z, a.z, b.z);
    opt_update(p.kind, p.hb, coef, g.w, pv.w, a.w, b.w);
    s0[i] = a;
    if (s1) s1[i] = b;
    lp[i] = pv;
  }
  if (tr) { p.trace[1] = p.trace[2] = p.trace[3] = globaltimer_ns(); }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(p.ticket, 1u);
    if (tk == (unsigned int)G - 1u) { *p.ticket = 0u; *p.epoch_ctr = epoch; if (p.advance) opt_advance_state(p.hb, p.kind); __threadfence(); }
  }
}

// NVLS variant of the two-shot bucket (multi-process): every rank's gradient bucket is bound at the same offset of ONE multicast
// object (nvls_drv.cuh), so rank r obtains the SUM of its slice with multimem.ld_reduce -- the addition happens inside the NVSwitch
// and only 1/W of the bucket crosses this rank's links on the way in (the push kernel sends (W-1)/W of it out).  Then x 1/W,
// optimizer on the slice, new parameters stored to every peer's parameter buffer (IPC-mapped arenas, as in the push kernel).
//   barrier (all ranks' gradients final) -> reduce + update + parameter push -> barrier (nobody still reads my gradients, all
//   parameter slices have landed) -> clear my gradients.
// The switch adds in a fixed order and every element is reduced exactly once (by its owner), so replicas stay bit-identical.
__device__ __forceinline__ float4 mm_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}

struct NvlsArParams {
  CommDev comm;                 // .flags = barrier flags (IPC arenas), .params[q] = parameter bucket of rank q, .grads unused
  const float* mc_grads;        // multicast address of the bucket's gradients
  float* grads_l;               // own (unicast) gradients of the bucket
  float* params_l; float* slot0; float* slot1;
  int n4, S4, kind, advance;
  HyperBlock* hb;
  unsigned int* epoch_ctr; unsigned int* ticket;
};

__global__ void __launch_bounds__(CIB_PAR_THREADS, 8) nvls_allreduce_opt_kernel(const NvlsArParams p) {
  const CommDev& c = p.comm;
  const int W = c.world, R = c.rank, G = gridDim.x, j = blockIdx.x;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  peer_barrier(c, 0, epoch);
  const OptCoef coef = opt_coefficients(p.hb, p.kind);
  const float inv_w = 1.0f / (float)W;
  const int base = R * p.S4;
  const int len = max(0, min(p.n4, base + p.S4) - base);
  const int per = (len + G - 1) / G;
  const int lo = min(len, j * per), hi = min(len, lo + per);
  float4* lp = reinterpret_cast<float4*>(p.params_l);
  float4* s0 = reinterpret_cast<float4*>(p.slot0);
  float4* s1 = p.slot1 ? reinterpret_cast<float4*>(p.slot1) : nullptr;
  for (int i = lo + threadIdx.x; i < hi; i += CIB_PAR_THREADS) {
    float4 g = mm_ld_reduce_f4(p.mc_grads + 4 * (size_t)(base + i));
    float4 pv = lp[base + i];
    float4 a = s0[base + i];
    float4 b = s1 ? s1[base + i] : zero4;
    g.x *= inv_w; g.y *= inv_w; g.z *= inv_w; g.w *= inv_w;
    opt_update(p.kind, p.hb, coef, g.x, pv.x, a.x, b.x);
    opt_update(p.kind, p.hb, coef, g.y, pv.y, a.y, b.y);
    opt_update(p.kind, p.hb, coef, g.z, pv.z, a.z, b.z);
    opt_update(p.kind, p.hb, coef, g.w, pv.w, a.w, b.w);
    s0[base + i] = a;
    if (s1) s1[base + i] = b;
    lp[base + i] = pv;
#pragma unroll 1
    for (int d = 1; d < W; ++d) st_peer_f4(c.params[(R + d) % W] + 4 * (size_t)(base + i), pv);
  }
  peer_barrier(c, 1, epoch);
  // everybody has read what it needed of my gradients: clear the part of the bucket this CTA index covers on every slice
  float4* lg = reinterpret_cast<float4*>(p.grads_l);
  const int per_all = (p.n4 + G - 1) / G;
  for (int i = j * per_all + threadIdx.x; i < min(p.n4, (j + 1) * per_all); i += CIB_PAR_THREADS) lg[i] = zero4;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(p.ticket, 1u);
    if (tk == (unsigned int)G - 1u) { *p.ticket = 0u; *p.epoch_ctr = epoch; if (p.advance) opt_advance_state(p.hb, p.kind); __threadfence(); }
  }
}

__global__ void stamp_kernel(unsigned long long* slot) { *slot = globaltimer_ns(); }

// Reassembles the optimizer moments that two-shot updates leave sharded (rank r only ever updates slice r of slot0 / slot1):
// every rank pulls slice q of both slot buffers from rank q.  Called by the host on ALL ranks before the slots are read
// (checkpoint, engine re-creation) or broadcast.  The slot buffers live in the peer-mapped arena.
struct GatherSlotsParams {
  CommDev comm;                        // .grads[q] = slot0 bucket of rank q, .params[q] = slot1 bucket of rank q (or null)
  int n4, S4;
  unsigned int* epoch_ctr; unsigned int* ticket;
};
__global__ void __launch_bounds__(512) gather_slots_kernel(const GatherSlotsParams p) {
  const CommDev& c = p.comm;
  const unsigned int epoch = *p.epoch_ctr + 1u;
  peer_barrier(c, 0, epoch);
  for (int q = 0; q < c.world; ++q) {
    if (q == c.rank) continue;
    const int lo = q * p.S4, hi = min(p.n4, lo + p.S4);
    for (int i = lo + blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += gridDim.x * blockDim.x) {
      reinterpret_cast<float4*>(c.grads[c.rank])[i] = ld_peer_f4(c.grads[q] + 4 * (size_t)i);
      if (c.params[0]) reinterpret_cast<float4*>(c.params[c.rank])[i] = ld_peer_f4(c.params[q] + 4 * (size_t)i);
    }
  }
  peer_barrier(c, 1, epoch);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(p.ticket, 1u);
    if (tk == gridDim.x - 1) { *p.ticket = 0; *p.epoch_ctr = epoch; __threadfence(); }
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
