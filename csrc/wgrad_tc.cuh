// 3x3 convolution weight gradient on tcgen05: both operands MN-major, the GEMM K dimension is "pixels".
//
//   dW[tap][ci][co] = sum_pix  X[pix + tap_offset][ci] * dZ[pix][co]          db[co] = sum_pix dZ[pix][co]
//
// expressed per tap as D_tap[M = co][N = ci] += A[K = 32 pixels][M]^T * B_tap[K = 32 pixels][N], where A is a
// (16 w x 2 h)-pixel box of dZ and B_tap the same box of X shifted by the tap offset -- both fetched by TMA from
// the NHWC tensors exactly as they sit in memory (channels contiguous = MN-major, no transposes, no im2col buffer).
// The bias gradient rides along as one more N = 32 accumulator whose B operand is a tile of ones.
// MN-major tf32 operands use the 128-byte swizzle with 32-byte atoms (TMA SWIZZLE_128B_ATOM_32B).
// All taps of a group accumulate side by side in TMEM (<= 512 columns) across the CTA's whole pixel slice; the
// epilogue then adds the CTA's partial sums into the flat gradient buffer with coalesced red.global.add.
//
// Replaces TensorFlow's Conv2DBackpropFilter for Keras Conv2D (reference rpv.py:46-48 via model.fit, rpv.py:103).
#pragma once
#include "ptx.cuh"

namespace cib {

struct WgradTcParams {
  CUtensorMap tmA;     // dZ : dims (CO_pad, Wo, Ho, N), box (CBa, 16, 2, 1)
  CUtensorMap tmB;     // X  : dims (CI_pad, Wi, Hi, N), box (CBb, 16, 2, 1)
  int CO, CI;          // channel counts as seen by the MMA: multiples of 32 (MN-major tf32 needs 128-byte rows)
  int ci_out;          // channels of X that really exist (= CI, or 16 when a 16-channel X is read through an
                       // overlapping tensor map whose 32-float rows span two neighbouring pixels; only the first
                       // 16 columns of every accumulator are then meaningful)
  int n_img, Ho, Wo;   // dZ pixel grid
  int chunks_w, chunks_h, num_chunks;     // (16 x 2)-pixel boxes
  int pad;             // 1 = same, 0 = valid
  int xstride;         // 2: strided convolution -- X is read through a tensor map with traversal stride 2 at (2 w0 + kw - pad, 2 h0 + kh - pad)
  int taps_per_group;  // 9, 3 or 1
  int tap_groups;      // 1, 3 or 9
  int m_tiles;         // ceil(CO / 128)
  int n_slices;        // pixel slices; grid = n_slices * tap_groups * m_tiles
  int stages;
  float* dW;           // [9][CI][CO] (Keras HWIO, padded)
  float* db;           // [CO]
};

__global__ void __launch_bounds__(192, 1) wgrad_tc_kernel(const __grid_constant__ WgradTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[8];
  __shared__ __align__(8) uint64_t empty_bar[8];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int S = p.stages;
  const int CO = p.CO, CI = p.CI;
  const int TPG = p.taps_per_group;

  int t = blockIdx.x;
  const int m_t = t % p.m_tiles; t /= p.m_tiles;
  const int tg = t % p.tap_groups;
  const int slice = t / p.tap_groups;
  const bool with_bias = (tg == 0);
  const int co0 = m_t * 128;
  const int co_cnt = min(128, CO - co0);                    // channels of dZ this CTA owns

  const int per = (p.num_chunks + p.n_slices - 1) / p.n_slices;
  const int ch_begin = slice * per;
  const int ch_end = min(p.num_chunks, ch_begin + per);
  const int nch = ch_end - ch_begin;

  const uint32_t rowb_a = 128u, rowb_b = 128u;
  const int a_blocks = co_cnt / 32;
  const int b_blocks = CI / 32;
  const uint32_t a_bytes = a_blocks * 32u * rowb_a;
  const uint32_t b_tap_bytes = b_blocks * 32u * rowb_b;
  const uint32_t b_bytes = TPG * b_tap_bytes;
  // Keep at least 16 KB after the A tile inside a stage: with M = 128 the hardware also reads the (ignored)
  // channel blocks past co_cnt, which must still be mapped shared memory.
  const uint32_t stage_bytes = a_bytes + (b_bytes < 16384u ? 16384u : b_bytes);
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ones_base = smem_base + S * stage_bytes;   // 8 rows x 128 B of 1.0f

  const int acc_cols = TPG * CI + (with_bias ? 32 : 0);
  uint32_t tmem_cols = 32; while ((int)tmem_cols < acc_cols) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA); tma_prefetch_desc(&p.tmB);
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full_bar[s]), 1); mbar_init(smem_u32(&empty_bar[s]), 1); }
    mbar_init(smem_u32(&tmem_full_bar), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  if (warp >= 2) {
    // ones tile for the bias-gradient MMA (256 floats), written through the generic proxy
    float* ones = reinterpret_cast<float*>(smem_raw + (ones_base - smem_u32(smem_raw)));
    const int tid = threadIdx.x - 64;
    ones[tid] = 1.0f; ones[tid + 128] = 1.0f;
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (nch > 0) {
    if (warp == 0) {
      if (lane == 0) {
        for (int i = 0; i < nch; ++i) {
          const int s = i % S; const uint32_t ph = (i / S) & 1;
          int c = ch_begin + i;
          const int cw = c % p.chunks_w; c /= p.chunks_w;
          const int chh = c % p.chunks_h; const int n = c / p.chunks_h;
          const int w0 = cw * 16, h0 = chh * 2;
          mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1, 0x100 | s);
          const uint32_t fb = smem_u32(&full_bar[s]);
          mbar_expect_tx(fb, a_bytes + b_bytes);
          const uint32_t a_dst = smem_base + s * stage_bytes;
          const uint32_t b_dst = a_dst + a_bytes;
          for (int b = 0; b < a_blocks; ++b)
            tma_load_4d(a_dst + b * (32u * rowb_a), &p.tmA, fb, co0 + b * 32, w0, h0, n);
          for (int tp = 0; tp < TPG; ++tp) {
            const int tap = tg * TPG + tp;
            const int kh = tap / 3, kw = tap % 3;
            for (int b = 0; b < b_blocks; ++b)
              tma_load_4d(b_dst + tp * b_tap_bytes + b * (32u * rowb_b), &p.tmB, fb,
                          b * 32, p.xstride * w0 + kw - p.pad, p.xstride * h0 + kh - p.pad, n);
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        const uint32_t idesc = make_idesc_tf32(128, CI, 1, 1);
        const uint32_t idesc_bias = make_idesc_tf32(128, 32, 1, 1);
        for (int i = 0; i < nch; ++i) {
          const int s = i % S; const uint32_t ph = (i / S) & 1;
          mbar_wait(smem_u32(&full_bar[s]), ph, 0x200 | s);
          tc_fence_after();
          const uint32_t a_src = smem_base + s * stage_bytes;
          const uint32_t b_src = a_src + a_bytes;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {                      // 4 x 8 pixel rows
            const uint64_t ad = make_sdesc(a_src + ks * 8 * rowb_a, 32u * rowb_a, 512, SWZ_128_B32);
            for (int tp = 0; tp < TPG; ++tp) {
              const uint64_t bd = make_sdesc(b_src + tp * b_tap_bytes + ks * 8 * rowb_b, 32u * rowb_b, 512, SWZ_128_B32);
              umma_tf32(tmem_base + tp * CI, ad, bd, idesc, (i > 0 || ks > 0) ? 1u : 0u);
            }
            if (with_bias) {
              const uint64_t od = make_sdesc(ones_base, 1024, 512, SWZ_128_B32);
              umma_tf32(tmem_base + TPG * CI, ad, od, idesc_bias, (i > 0 || ks > 0) ? 1u : 0u);
            }
          }
          umma_commit(smem_u32(&empty_bar[s]));
        }
        umma_commit(smem_u32(&tmem_full_bar));
      }
    } else {
      const int q = warp & 3;
      const int co = co0 + q * 32 + lane;
      mbar_wait(smem_u32(&tmem_full_bar), 0, 0x300);
      tc_fence_after();
      const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
      const bool row_ok = (q * 32 + lane) < co_cnt;
      for (int tp = 0; tp < TPG; ++tp) {
        const int tap = tg * TPG + tp;
        for (int c0 = 0; c0 < p.ci_out; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(tlane + tp * CI + c0, v);
          tmem_ld_wait();
          if (row_ok) {
            float* dst = p.dW + ((size_t)tap * p.ci_out + c0) * CO + co;
#pragma unroll
            for (int j = 0; j < 16; ++j) red_add_f32(dst + (size_t)j * CO, __uint_as_float(v[j]));
          }
        }
      }
      if (with_bias) {
        uint32_t v[16];
        tmem_ld16(tlane + TPG * CI, v);
        tmem_ld_wait();
        if (row_ok) red_add_f32(p.db + co, __uint_as_float(v[0]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace cib
