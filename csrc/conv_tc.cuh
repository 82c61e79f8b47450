// 3x3 convolution forward and data-gradient as an implicit GEMM on tcgen05 (kind::tf32, fp32 in TMEM).
//
// For every filter tap the A operand is a 128-pixel x KK-channel tile of the NHWC source tensor, fetched by TMA
// straight from global memory at the tap's spatial offset (tiled-mode tensor map; out-of-bounds coordinates are
// zero-filled by the TMA unit, which is exactly "same" padding).  The B operand is that tap's KK x NN filter
// slice, fetched from the Keras-layout (kh,kw,Cin,Cout) kernel without any transposition:
//   forward : K = Cin,  N = Cout  -> slice is [K rows][N contiguous]  = MN-major B
//   dgrad   : K = Cout, N = Cin   -> slice is [N rows][K contiguous]  = K-major  B
// 9 taps x (KK/8) tcgen05.mma accumulate one 128 x NN tile in TMEM; two TMEM accumulator stages let the
// epilogue of tile i overlap the MMAs of tile i+1.  Persistent CTAs (one per SM) walk the tile list.
//
// A 128-row tile is 4 TMA boxes of (16 w x 2 h) pixels, so a 2x2 pooling window always lives inside one warp's
// 32 TMEM lanes and the fused bias+ReLU+max-pool epilogue needs only warp shuffles.
//
// Stored activations carry the pooling arg-max position in the 2 mantissa LSBs (tf32 MMAs ignore the low 13
// mantissa bits); "value != 0" doubles as the ReLU/dropout mask for the backward pass.
//
// Replaces Keras Conv2D+MaxPooling2D forward and autograd dgrad (reference rpv.py:44-50, mnist.py:49-52).
#pragma once
#include "ptx.cuh"

namespace cib {

enum { CONV_EPI_FWD = 0, CONV_EPI_DGRAD_UNPOOL = 1, CONV_EPI_DGRAD_MASK = 2 };

// Device-resident hyper-parameter / state block shared by all kernels of one model (graph-replay safe).
struct HyperBlock {
  float lr, beta1, beta2, eps, rho, momentum, schedule_decay, decay;
  float m_schedule[2];        // Nadam running product, double buffered by iteration parity
  unsigned int iterations;    // optimizer steps done so far
  unsigned int seed;
  unsigned int ticket;        // last-block election for the optimizer kernel
  unsigned int training;      // 1 while dropout is active
  float loss_sum, correct_sum, sample_sum;   // epoch accumulators
  float step_loss, step_acc;                 // last step's batch loss / accuracy
  unsigned int cursor_wrap;  // 1: the batch cursor restarts at 0 when the next batch would run past the dataset (steady-state timing loops)
  // ---- dataset block (written as one 32-byte copy by the host, see push_dataset) ----
  unsigned int batch_cursor;  // position in the device-resident permutation
  int n_total;                // rows of the current device-resident dataset
  const float* data_x;        // current dataset (switched between train / validation without re-capturing graphs)
  const float* data_y;
  const int* perm;            // nullptr = identity order
};

__device__ __forceinline__ void advance_batch_cursor(HyperBlock* hb, int B) {
  unsigned int c = hb->batch_cursor + (unsigned int)B;
  if (hb->cursor_wrap && c + (unsigned int)B > (unsigned int)hb->n_total) c = 0u;
  hb->batch_cursor = c;
}

struct ConvTcParams {
  CUtensorMap tmA;     // source tensor, dims (Ca, Wa, Ha, N), box (CBa, 16, 2, 1)
  CUtensorMap tmB;     // filter, dims (Cout_pad, 9*Cin_pad)
  int b_mn;            // 1 = forward (B MN-major), 0 = dgrad (B K-major)
  int KK, NN;          // GEMM K (source channels) and N (result channels), padded: 16 or a multiple of 32
  int cin_pad;         // rows per tap in the filter matrix
  int n_img, Ho, Wo;   // result pixel grid (before pooling)
  int TW;              // tile width in pixels: 16 or 32 (tile height = 128 / TW)
  int tiles_w, tiles_h, num_tiles;
  int ntaps;           // 9
  int tap_dh[9], tap_dw[9], tap_row[9];   // source offset per tap and filter-matrix row block
  int stages;
  int sstride;         // source stride: 2 = strided forward convolution (the tensor map's traversal stride does the skipping), else 1
  // Strided dgrad = four unit-stride convolutions, one per parity class (a, b) of the input pixel: class pixels (2i + a, 2j + b)
  // receive only the taps with kh = a + pad, kw = b + pad (mod 2).  The result of a class is a sub-grid that is scattered with
  // stride `ostride` and offset (oa, ob) into the full grid [Hfull][Wfull] (also the grid of act_prev).  ostride = 1: plain.
  int ostride, oa, ob, Hfull, Wfull;
  int kchunk;          // 1: a pipeline stage is ONE 32-channel K chunk of a tap (wide layers whose whole-tap stage does not fit twice)
  int epi;
  // forward epilogue
  const float* bias;
  float* out;          // [n][Hs][Ws][NN] with Hs = pool ? Ho/2 : Ho
  int pool;
  float drop_rate;     // 0 = no dropout
  unsigned int layer_id;
  const HyperBlock* hb;
  // dgrad epilogues
  const float* act_prev;   // stored activation on the result grid [n][Ho][Wo][NN]
  float* dz_out;           // UNPOOL: [n][Hf][Wf][NN]; MASK: [n][Ho][Wo][NN]
  int Hf, Wf;
  float scale;
};

__device__ __forceinline__ void conv_tile_coords(const ConvTcParams& p, int tile, int& n, int& h0, int& w0) {
  const int tw = tile % p.tiles_w; tile /= p.tiles_w;
  const int th = tile % p.tiles_h; n = tile / p.tiles_h;
  w0 = tw * p.TW; h0 = th * (128 / p.TW);
}

__global__ void __launch_bounds__(192, 1) conv_tc_kernel(const __grid_constant__ ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[8];
  __shared__ __align__(8) uint64_t empty_bar[8];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int S = p.stages;
  const int KK = p.KK, NN = p.NN;
  // K-chunked mode (wide layers, e.g. 128 -> 256 channels: one tap is 64 KB of A + 128 KB of filter): a stage holds one 32-channel
  // chunk of one tap (16 KB + 32 x NN x 4 B), so four or more stages fit and TMA keeps running ahead of the MMAs
  const bool kch = p.kchunk != 0;
  const int kcs = kch ? KK / 32 : 1;
  const uint32_t a_bytes = kch ? 128u * 128u : 128u * KK * 4u;
  const uint32_t b_bytes = kch ? 32u * NN * 4u : (uint32_t)KK * NN * 4u;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t rowb_a = KK >= 32 ? 128u : 64u;           // bytes per A row chunk
  const int a_chunks = KK >= 32 ? KK / 32 : 1;
  const int cb_a = KK >= 32 ? 32 : 16;

  uint32_t tmem_cols = 32; while ((int)tmem_cols < 2 * NN) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA); tma_prefetch_desc(&p.tmB);
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full_bar[s]), 1); mbar_init(smem_u32(&empty_bar[s]), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(smem_u32(&tfull_bar[s]), 1); mbar_init(smem_u32(&tempty_bar[s]), 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        int n, h0, w0; conv_tile_coords(p, tile, n, h0, w0);
        for (int tapk = 0; tapk < p.ntaps * kcs; ++tapk, ++it) {
          const int tap = tapk / kcs, kq = tapk - tap * kcs;
          const int s = it % S; const uint32_t ph = (it / S) & 1;
          mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1, 0x100 | s);
          const uint32_t fb = smem_u32(&full_bar[s]);
          mbar_expect_tx(fb, stage_bytes);
          const uint32_t a_dst = smem_base + s * stage_bytes;
          const uint32_t b_dst = a_dst + a_bytes;
          if (kch) {
#pragma unroll
            for (int sb = 0; sb < 4; ++sb) {
              const int dw = p.TW == 32 ? (sb & 1) * 16 : 0;
              const int dh = p.TW == 32 ? (sb >> 1) * 2 : sb * 2;
              tma_load_4d(a_dst + sb * (32u * 128u), &p.tmA, fb, kq * 32, p.sstride * (w0 + dw) + p.tap_dw[tap], p.sstride * (h0 + dh) + p.tap_dh[tap], n);
            }
            if (p.b_mn) { for (int b = 0; b < NN / 32; ++b) tma_load_2d(b_dst + b * 4096u, &p.tmB, fb, b * 32, p.tap_row[tap] + kq * 32); }   // box (32 n, 32 k)
            else        tma_load_2d(b_dst, &p.tmB, fb, kq * 32, p.tap_row[tap]);                                                                // box (32 k, NN n)
            continue;
          }
          for (int kc = 0; kc < a_chunks; ++kc) {
#pragma unroll
            for (int sb = 0; sb < 4; ++sb) {
              const int dw = p.TW == 32 ? (sb & 1) * 16 : 0;
              const int dh = p.TW == 32 ? (sb >> 1) * 2 : sb * 2;
              tma_load_4d(a_dst + kc * (128u * rowb_a) + sb * (32u * rowb_a), &p.tmA, fb,
                          kc * cb_a, p.sstride * (w0 + dw) + p.tap_dw[tap], p.sstride * (h0 + dh) + p.tap_dh[tap], n);
            }
          }
          if (p.b_mn) {
            // [K = KK rows][NN contiguous], one box per 32-wide column block (NN is a multiple of 32)
            for (int b = 0; b < NN / 32; ++b)
              tma_load_2d(b_dst + b * (KK * 128u), &p.tmB, fb, b * 32, p.tap_row[tap]);
          } else {
            // [N = NN rows][KK contiguous], one box per 32-wide K chunk
            for (int kc = 0; kc < a_chunks; ++kc)
              tma_load_2d(b_dst + kc * (NN * rowb_a), &p.tmB, fb, kc * cb_a, p.tap_row[tap]);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(128, NN, 0, p.b_mn ? 1 : 0);
      const uint32_t a_layout = KK >= 32 ? SWZ_128 : SWZ_64;
      const uint32_t a_sbo = 8 * rowb_a;
      int it = 0, ti = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++ti) {
        const int as = ti & 1; const uint32_t aph = (ti >> 1) & 1;
        mbar_wait(smem_u32(&tempty_bar[as]), aph ^ 1, 0x400 | as);
        tc_fence_after();
        const uint32_t tacc = tmem_base + as * NN;
        for (int tapk = 0; tapk < p.ntaps * kcs; ++tapk, ++it) {
          const int tap = tapk / kcs;
          const int s = it % S; const uint32_t ph = (it / S) & 1;
          mbar_wait(smem_u32(&full_bar[s]), ph, 0x200 | s);
          tc_fence_after();
          const uint32_t a_src = smem_base + s * stage_bytes;
          const uint32_t b_src = a_src + a_bytes;
          if (kch) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              const uint64_t ad = make_sdesc(a_src + ks * 32, 16, a_sbo, SWZ_128);
              const uint64_t bd = p.b_mn ? make_sdesc(b_src + ks * 1024u, 4096u, 512, SWZ_128_B32) : make_sdesc(b_src + ks * 32, 16, 1024, SWZ_128);
              umma_tf32(tacc, ad, bd, idesc, (tapk > 0 || ks > 0) ? 1u : 0u);
            }
            umma_commit(smem_u32(&empty_bar[s]));
            continue;
          }
          for (int ks = 0; ks < KK / 8; ++ks) {
            uint64_t ad, bd;
            if (KK >= 32) ad = make_sdesc(a_src + (ks >> 2) * (128u * 128u) + (ks & 3) * 32, 16, a_sbo, a_layout);
            else          ad = make_sdesc(a_src + ks * 32, 16, a_sbo, a_layout);
            if (p.b_mn) {
              bd = make_sdesc(b_src + ks * 1024u, KK * 128u, 512, SWZ_128_B32);   // MN-major tf32: 4-row atoms
            } else {
              if (KK >= 32) bd = make_sdesc(b_src + (ks >> 2) * (NN * 128u) + (ks & 3) * 32, 16, 1024, SWZ_128);
              else          bd = make_sdesc(b_src + ks * 32, 16, 512, SWZ_64);
            }
            umma_tf32(tacc, ad, bd, idesc, (tap > 0 || ks > 0) ? 1u : 0u);
          }
          umma_commit(smem_u32(&empty_bar[s]));
        }
        umma_commit(smem_u32(&tfull_bar[as]));
      }
    }
  } else {
    // =============================== epilogue warps ===============================
    const int q = warp & 3;
    const int dw = p.TW == 32 ? (q & 1) * 16 : 0;
    const int dh = p.TW == 32 ? (q >> 1) * 2 : q * 2;
    const int lh = lane >> 4, lw = lane & 15;
    const unsigned int step = p.hb ? p.hb->iterations : 0u;
    const unsigned int seed = p.hb ? p.hb->seed : 0u;
    const bool do_drop = p.drop_rate > 0.f && p.hb && p.hb->training;
    const float keep_scale = do_drop ? 1.0f / (1.0f - p.drop_rate) : 1.0f;
    int ti = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++ti) {
      const int as = ti & 1; const uint32_t aph = (ti >> 1) & 1;
      int n, h0, w0; conv_tile_coords(p, tile, n, h0, w0);
      const int h = h0 + dh + lh, w = w0 + dw + lw;
      const bool inb = h < p.Ho && w < p.Wo;
      mbar_wait(smem_u32(&tfull_bar[as]), aph, 0x300 | as);
      tc_fence_after();
      const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16) + as * NN;
      for (int c0 = 0; c0 < NN; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tlane + c0, v);
        tmem_ld_wait();
        if (p.epi == CONV_EPI_FWD) {
          float x[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) x[j] = fmaxf(__uint_as_float(v[j]) + __ldg(p.bias + c0 + j), 0.f);
          if (p.pool) {
            const int Hs = p.Ho >> 1, Ws = p.Wo >> 1;
            const int gi = (lane & 1) + 2 * lh;          // my position inside the 2x2 window
            uint32_t pos[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              float xv = x[j]; uint32_t pv = gi;
              float xo = __shfl_xor_sync(0xffffffffu, xv, 1); uint32_t po = __shfl_xor_sync(0xffffffffu, pv, 1);
              if (xo > xv || (xo == xv && po < pv)) { xv = xo; pv = po; }
              xo = __shfl_xor_sync(0xffffffffu, xv, 16); po = __shfl_xor_sync(0xffffffffu, pv, 16);
              if (xo > xv || (xo == xv && po < pv)) { xv = xo; pv = po; }
              x[j] = xv; pos[j] = pv;
            }
            const int phh = h >> 1, pww = w >> 1;
            if (phh < Hs && pww < Ws) {
              // the 4 lanes of a window each store 4 of the 16 channels
              const size_t oidx = (((size_t)n * Hs + phh) * Ws + pww) * NN + c0 + 4 * gi;
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float xv = 0.f; uint32_t pv = 0;
#pragma unroll
                for (int j = 0; j < 16; ++j) if (j == 4 * gi + e) { xv = x[j]; pv = pos[j]; }
                if (do_drop) {
                  const float u = dropout_uniform(seed, step, p.layer_id, (uint32_t)(oidx + e));
                  xv = u >= p.drop_rate ? xv * keep_scale : 0.f;
                }
                o[e] = (__float_as_uint(xv) & ~3u) | pv;
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
                float xv = x[4 * j + e];
                if (do_drop) {
                  const float u = dropout_uniform(seed, step, p.layer_id, (uint32_t)(oidx + 4 * j + e));
                  xv = u >= p.drop_rate ? xv * keep_scale : 0.f;
                }
                y[e] = __uint_as_float(__float_as_uint(xv) & ~3u);
              }
              dst[j] = make_float4(y[0], y[1], y[2], y[3]);
            }
          }
        } else if (inb) {
          const int OH = h * p.ostride + p.oa, OW = w * p.ostride + p.ob;       // position in the full grid (strided dgrad classes)
          const float* a = p.act_prev + (((size_t)n * p.Hfull + OH) * p.Wfull + OW) * NN + c0;
          float g[16]; int pos[16];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint4 ab = *reinterpret_cast<const uint4*>(a + 4 * j);
            const uint32_t bits[4] = {ab.x, ab.y, ab.z, ab.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              pos[4 * j + e] = bits[e] & 3;
              g[4 * j + e] = (bits[e] & ~3u) ? __uint_as_float(v[4 * j + e]) * p.scale : 0.0f;
            }
          }
          if (p.epi == CONV_EPI_DGRAD_MASK) {
            float4* dst = reinterpret_cast<float4*>(p.dz_out + (((size_t)n * p.Hfull + OH) * p.Wfull + OW) * NN + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j) dst[j] = make_float4(g[4 * j], g[4 * j + 1], g[4 * j + 2], g[4 * j + 3]);
          } else {
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
              const int hh = 2 * h + (qq >> 1), ww = 2 * w + (qq & 1);
              float4* dst = reinterpret_cast<float4*>(p.dz_out + (((size_t)n * p.Hf + hh) * p.Wf + ww) * NN + c0);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                dst[j] = make_float4(pos[4 * j] == qq ? g[4 * j] : 0.f, pos[4 * j + 1] == qq ? g[4 * j + 1] : 0.f,
                                     pos[4 * j + 2] == qq ? g[4 * j + 2] : 0.f, pos[4 * j + 3] == qq ? g[4 * j + 3] : 0.f);
            }
          }
        }
      }
      // release this accumulator stage back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[as]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace cib
