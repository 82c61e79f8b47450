// First convolution layer (1 or 3 input channels) on the tensor cores.
//
// With K = 9*Cin = 9 or 27 the layer is far too thin for a per-tap implicit GEMM, and the CUDA-core kernels of
// kernels_basic.cuh are instruction-bound (ncu: 14 M / 19 M warp instructions for forward / wgrad of the 64x64x3 RPV
// layer, 30 / 41 us).  Here the loader warps materialise the im2col matrix of a (16 w x 8 h)-pixel tile in shared memory
// and the layer becomes one small GEMM per tile.
//
// Producer, in two steps (measured, B200):
//   1. the (10 x 18)-pixel halo of the tile is staged once with 4-byte cp.async (zero-fill outside the image), two tiles
//      ahead of its use;
//   2. the eight loader warps expand it shared->shared into the swizzled operand tile with LDS/STS.
//   Copying every im2col element straight from global memory with its own cp.async was 2x SLOWER than the CUDA-core
//   kernel: a warp-level LDGSTS retires only every ~15 cycles per SM whatever its width or access pattern (44..51 us
//   forward with the copies, 21 us without), so the 9x redundancy of im2col must not go through that path.
//
//   forward  D[128 px][CO] = A[128 px][32] x W^T[CO][32], both K-major SWIZZLE_128B: 4 MMAs per tile; four TMEM accumulator
//            stages (with two, the MMA -> epilogue -> MMA round trip paced the kernel: 0.55 us per tile with all work
//            removed); epilogue = bias + ReLU + 2x2 max-pool with arg-max bits + dropout (as conv_tc2).
//   wgrad    (conv1_tc_wgrad.cuh) D[taps][CO] += A^T[32 taps][128 px] x dZ^T[CO][128 px], both K-major (K = pixels).
//
// Reference call sites of the op: rpv.py:46-49 (first Conv2D + MaxPooling2D), mnist.py:49 (first Conv2D, no pool).
#pragma once
#include "ptx.cuh"
#include "conv_tc2.cuh"

namespace cib {

struct Conv1TcParams {
  const float* x;        // [n][Hi][Wi][CI], CI = 1 or 3
  const float* w;        // [9*CI][CO]
  const float* bias; float* out;
  int n_img, Hi, Wi, CI, CO, Ho, Wo, pad, pool;
  int tiles_w, tiles_h, num_tiles, stages;
  float drop_rate; unsigned int layer_id; const HyperBlock* hb;
  // wgrad: g = gradient at the stored grid (pooled if pool), already masked; act = stored activation (arg-max bits)
  const float* g; const float* act; float* dW; float* db;
  int debug;             // experiments: 1 skip the halo copies, 2 skip the epilogue math, 4 skip the dZ rebuild, 8 skip the expansion
};

static constexpr int C1_FWD_THREADS = 576;     // warp 0 MMA, 1 spare, 2..9 epilogue, 10..17 loaders
static constexpr int C1_WG_THREADS = 448;      // warp 0 MMA, 1 spare, 2..5 epilogue, 6..13 loaders
static constexpr int C1_LOADERS = 256;
static constexpr uint32_t C1_TILE_BYTES = 128u * 128u;
static constexpr int C1_HALO_SLOTS = 4;        // ring of staged halos: prefetch distance 2 + the two tiles being expanded
static constexpr int C1_HALO_AHEAD = 2;
static constexpr int C1_HALO_FLOATS = 10 * 18 * 3 + 20;   // one slot (padded)
static constexpr int C1_ACC_STAGES = 4;

__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void loaders_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr)); return v;
}
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }

__device__ __forceinline__ void c1_tile_coords(const Conv1TcParams& p, int tile, int& n, int& h0, int& w0) {
  const int tw = tile % p.tiles_w; tile /= p.tiles_w;
  const int th = tile % p.tiles_h; n = tile / p.tiles_h;
  w0 = tw * 16; h0 = th * 8;
}

// ---- step 1: halo staging.  Thread lt copies elements e = lt + 256 j (j < 3) of the [10][18*CI] halo.
struct C1Halo { int row[3], col[3], px[3]; };
__device__ __forceinline__ void c1_make_halo(C1Halo& H, int lt, int CI) {
  const int rowlen = 18 * CI;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int e = lt + j * C1_LOADERS;
    H.row[j] = e / rowlen; H.col[j] = e - H.row[j] * rowlen; H.px[j] = H.col[j] / CI;
    if (e >= 10 * rowlen) H.row[j] = -1;
  }
}
__device__ __forceinline__ void c1_issue_halo(const C1Halo& H, const Conv1TcParams& p, int tile, uint32_t slot_base) {
  if (p.debug & 1) return;
  int n, h0, w0; c1_tile_coords(p, tile, n, h0, w0);
  const int gh0 = h0 - p.pad, gw0 = w0 - p.pad;
  const float* img = p.x + (size_t)n * p.Hi * p.Wi * p.CI;
  const int rowlen = 18 * p.CI;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    if (H.row[j] >= 0) {
      const int gh = gh0 + H.row[j], gw = gw0 + H.px[j];
      const bool ok = (unsigned)gh < (unsigned)p.Hi && (unsigned)gw < (unsigned)p.Wi;
      const float* src = ok ? img + ((long long)gh * p.Wi + gw0) * p.CI + H.col[j] : img;
      cp_async4(slot_base + (uint32_t)(H.row[j] * rowlen + H.col[j]) * 4u, src, ok ? 4u : 0u);
    }
  }
}

// =====================================================================================================================
// forward
// =====================================================================================================================
__global__ void __launch_bounds__(C1_FWD_THREADS, 1) conv1_tc_fwd_kernel(const Conv1TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[8];
  __shared__ __align__(8) uint64_t empty_bar[8];
  __shared__ __align__(8) uint64_t tfull_bar[C1_ACC_STAGES];
  __shared__ __align__(8) uint64_t tempty_bar[C1_ACC_STAGES];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float bias_s[64];
  __shared__ __align__(16) float halo_s[C1_HALO_SLOTS][C1_HALO_FLOATS];

  pdl_launch();
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int S = p.stages, NN = p.CO;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t b_bytes = (uint32_t)NN * 128u;                      // W^T: one 128-byte row (32 k values) per output channel
  const uint32_t a_base = smem_base + ((b_bytes + 1023u) & ~1023u);
  uint32_t tmem_cols = 32; while ((int)tmem_cols < C1_ACC_STAGES * NN) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full_bar[s]), C1_LOADERS / 32); mbar_init(smem_u32(&empty_bar[s]), 1); }
    for (int s = 0; s < C1_ACC_STAGES; ++s) { mbar_init(smem_u32(&tfull_bar[s]), 1); mbar_init(smem_u32(&tempty_bar[s]), 4); }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  for (int i = threadIdx.x; i < NN; i += blockDim.x) bias_s[i] = p.bias[i];
  {
    // zero the filter tile and the A stages (columns >= 9*CI are never written again)
    float4* z4 = reinterpret_cast<float4*>(smem_al);
    const uint32_t n4 = ((a_base - smem_base) + S * C1_TILE_BYTES) / 16;
    for (uint32_t i = threadIdx.x; i < n4; i += blockDim.x) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    // filter, K-major: element (co, k) at row co, 16-byte piece (k >> 2) ^ (co & 7)
    const int K = 9 * p.CI;
    for (int i = threadIdx.x; i < K * NN; i += blockDim.x) {
      const int k = i / NN, co = i - k * NN;
      *reinterpret_cast<float*>(smem_al + co * 128u + (((k >> 2) ^ (co & 7)) << 4) + (k & 3) * 4u) = p.w[i];
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_smem, 0);
  pdl_wait();                                                        // filter / bias above are parameters: constant within a step

  if (warp == 0) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_tf32(128, NN, 0, 0);
    const uint64_t a_desc0 = make_sdesc(a_base, 16, 1024, SWZ_128);
    const uint64_t b_desc0 = make_sdesc(smem_base, 16, 1024, SWZ_128);
    const uint32_t stage_u = C1_TILE_BYTES >> 4;
    // Tiles are issued four at a time, K step by K step across the four accumulator stages: back-to-back tcgen05.mma into the
    // SAME accumulator are serialised by its read-modify-write latency (~100 cycles for these tiny MMAs, measured).
    const int ntile = (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    for (int g0 = 0; g0 < ntile; g0 += C1_ACC_STAGES) {
      const int cnt = min(C1_ACC_STAGES, ntile - g0);
      const uint32_t aph = (g0 / C1_ACC_STAGES) & 1;
      for (int t = 0; t < cnt; ++t) {
        const int ti = g0 + t;
        mbar_wait(smem_u32(&tempty_bar[t]), aph ^ 1, 0x400 | t);
        mbar_wait(smem_u32(&full_bar[ti % S]), (ti / S) & 1, 0x200 | (ti % S));
      }
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < 4; ++ks)
        for (int t = 0; t < cnt; ++t)
          if (leader) umma_tf32(tmem_base + t * NN, a_desc0 + (uint64_t)(((g0 + t) % S) * stage_u) + 2u * ks, b_desc0 + 2u * ks, idesc, ks > 0 ? 1u : 0u);
      for (int t = 0; t < cnt; ++t) {
        if (leader) umma_commit(smem_u32(&empty_bar[(g0 + t) % S]));
        if (leader) umma_commit(smem_u32(&tfull_bar[t]));
      }
      __syncwarp();
    }
  } else if (warp >= 10) {
    // ====================== loaders: halo staging (async, 2 tiles ahead) + shared->shared im2col ======================
    const int lt = threadIdx.x - 320, lw = lt >> 5;
    C1Halo H; c1_make_halo(H, lt, p.CI);
    const int rowlen = 18 * p.CI;
    // expansion: one warp instruction = one pixel row (CI = 3: lane = column k) or two rows (CI = 1: 16 lanes per row)
    const int KL = p.CI == 1 ? 16 : 32, rpw = 32 / KL;
    const int k = lane % KL, rsub = lane / KL;
    const bool kact = k < 9 * p.CI;
    const int tap = k / p.CI, ci = k - tap * p.CI, kh = tap / 3, kw = tap - kh * 3;
    const uint32_t koff = (uint32_t)(kh * rowlen + kw * p.CI + ci) * 4u;
    const int row0 = lw * rpw + rsub, rstep = 8 * rpw, nrows = 128 / rstep;
    const uint32_t kpiece = k >> 2, kword = (k & 3) * 4u;
    const uint32_t halo_base = smem_u32(&halo_s[0][0]);
    const int ntile = (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;    // tiles of this CTA
    for (int i = 0; i < C1_HALO_AHEAD; ++i) {
      if (i < ntile) c1_issue_halo(H, p, blockIdx.x + i * gridDim.x, halo_base + (i % C1_HALO_SLOTS) * (C1_HALO_FLOATS * 4u));
      cp_async_commit();
    }
    for (int i = 0; i < ntile; ++i) {
      if (i + C1_HALO_AHEAD < ntile)
        c1_issue_halo(H, p, blockIdx.x + (i + C1_HALO_AHEAD) * gridDim.x, halo_base + ((i + C1_HALO_AHEAD) % C1_HALO_SLOTS) * (C1_HALO_FLOATS * 4u));
      cp_async_commit();                                    // (possibly empty) group: keeps the group count per tile fixed
      cp_async_wait<C1_HALO_AHEAD>();                       // this thread's part of halo i has landed
      loaders_bar();                                        // ... and everybody else's
      const int s = i % S; const uint32_t ph = (i / S) & 1;
      if (lane == 0) mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1, 0x100 | s);
      __syncwarp();
      if (kact && !(p.debug & 8)) {
        const uint32_t src0 = halo_base + (i % C1_HALO_SLOTS) * (C1_HALO_FLOATS * 4u) + koff;
        const uint32_t dst0 = a_base + s * C1_TILE_BYTES + kword;
#pragma unroll 4
        for (int j = 0; j < nrows; ++j) {
          const int m = row0 + j * rstep;
          const float v = lds_f32(src0 + (uint32_t)((m >> 4) * rowlen + (m & 15) * p.CI) * 4u);
          sts_f32(dst0 + m * 128u + ((kpiece ^ (uint32_t)(m & 7)) << 4), v);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&full_bar[s]));
    }
  } else if (warp >= 2) {
    // ====================== epilogue: warps 2..5 take even tiles, warps 6..9 odd tiles ======================
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int lh = lane >> 4, lw = lane & 15;
    const unsigned int step = p.hb ? p.hb->iterations : 0u;
    const unsigned int seed = p.hb ? p.hb->seed : 0u;
    const bool do_drop = p.drop_rate > 0.f && p.hb && p.hb->training;
    const float keep_scale = do_drop ? 1.0f / (1.0f - p.drop_rate) : 1.0f;
    const int gi = (lane & 1) + 2 * lh;
    int ti = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++ti) {
      if ((ti & 1) != half) continue;
      const int as = ti % C1_ACC_STAGES; const uint32_t aph = (ti / C1_ACC_STAGES) & 1;
      int n, h0, w0; c1_tile_coords(p, tile, n, h0, w0);
      const int h = h0 + 2 * q + lh, w = w0 + lw;
      const bool inb = h < p.Ho && w < p.Wo;
      mbar_wait(smem_u32(&tfull_bar[as]), aph, 0x300 | as);
      tc_fence_after();
      const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16) + as * NN;
      for (int c0 = 0; c0 < NN && !(p.debug & 2); c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tlane + c0, v);
        tmem_ld_wait();
        if (p.pool) {
          const int Hs2 = p.Ho >> 1, Ws2 = p.Wo >> 1;
          const int wbase = lane & ~17;
          float x[16], m[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) x[j] = inb ? fmaxf(__uint_as_float(v[j]) + bias_s[c0 + j], 0.f) : 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) m[j] = fmaxf(x[j], __shfl_xor_sync(0xffffffffu, x[j], 1));
#pragma unroll
          for (int j = 0; j < 16; ++j) m[j] = fmaxf(m[j], __shfl_xor_sync(0xffffffffu, m[j], 16));
          uint32_t eq[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) eq[j] = __ballot_sync(0xffffffffu, x[j] == m[j]) >> wbase;
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float mv = m[e]; uint32_t qv = eq[e];
            mv = gi == 1 ? m[4 + e] : mv;  qv = gi == 1 ? eq[4 + e] : qv;
            mv = gi == 2 ? m[8 + e] : mv;  qv = gi == 2 ? eq[8 + e] : qv;
            mv = gi == 3 ? m[12 + e] : mv; qv = gi == 3 ? eq[12 + e] : qv;
            const uint32_t w4 = (qv & 3u) | ((qv >> 14) & 12u);
            const uint32_t pv = (uint32_t)__ffs((int)w4) - 1u;
            o[e] = (__float_as_uint(mv) & ~3u) | (pv & 3u);
          }
          const int phh = h >> 1, pww = w >> 1;
          if (phh < Hs2 && pww < Ws2) {
            const size_t oidx = (((size_t)n * Hs2 + phh) * Ws2 + pww) * NN + c0 + 4 * gi;
            if (do_drop) {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float u = dropout_uniform(seed, step, p.layer_id, (uint32_t)(oidx + e));
                const float xv = __uint_as_float(o[e] & ~3u);
                o[e] = (u >= p.drop_rate ? (__float_as_uint(xv * keep_scale) & ~3u) : 0u) | (o[e] & 3u);
              }
            }
            *reinterpret_cast<uint4*>(p.out + oidx) = make_uint4(o[0], o[1], o[2], o[3]);
          }
        } else if (inb) {
          const size_t oidx = (((size_t)n * p.Ho + h) * p.Wo + w) * NN + c0;
          float4* dst = reinterpret_cast<float4*>(p.out + oidx);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float y[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float xv = fmaxf(__uint_as_float(v[4 * j + e]) + bias_s[c0 + 4 * j + e], 0.f);
              if (do_drop) {
                const float u = dropout_uniform(seed, step, p.layer_id, (uint32_t)(oidx + 4 * j + e));
                xv = u >= p.drop_rate ? xv * keep_scale : 0.f;
              }
              y[e] = __uint_as_float(__float_as_uint(xv) & ~3u);
            }
            dst[j] = make_float4(y[0], y[1], y[2], y[3]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[as]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace cib

#include "conv1_tc_wgrad.cuh"
