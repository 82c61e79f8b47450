// CUDA-core kernels: the first conv layer (Cin = 1 or 3 is too thin for an MMA K dimension), loss heads,
// the multi-tensor optimizers, the batch gather, and shape-agnostic fall-backs (strided convs, odd topologies)
// that share the buffer conventions of the tcgen05 kernels so either path can serve any layer.
#pragma once
#include "ptx.cuh"
#include "conv_tc.cuh"   // HyperBlock

namespace cib {

// =============================================================================================
// Conv forward, direct form.  One thread = one output site (a 2x2 window if pooled) x 8 output channels.
// Fuses bias + ReLU + 2x2 max-pool (+ arg-max bits) + dropout.  Weights: [9][CI][CO] (Keras HWIO, CO padded).
// Reference: Conv2D/MaxPooling2D/Dropout at rpv.py:46-50, mnist.py:49-52.
// =============================================================================================
struct ConvDirectParams {
  const float* x;      // [n][Hi][Wi][CI]
  const float* w;      // [KH*KW][CI][CO]
  const float* bias;   // [CO]
  float* out;          // [n][Hs][Ws][CO]
  int n_img, Hi, Wi, CI, CO, Ho, Wo, KH, KW, stride, pad, pool;
  float drop_rate; unsigned int layer_id;
  const HyperBlock* hb;
  int w_in_smem;
};

__global__ void __launch_bounds__(256) conv_fwd_direct_kernel(const ConvDirectParams p) {
  extern __shared__ float wsm[];
  const int wcount = p.KH * p.KW * p.CI * p.CO;
  if (p.w_in_smem) {
    for (int i = threadIdx.x * 4; i < wcount; i += blockDim.x * 4)
      *reinterpret_cast<float4*>(wsm + i) = *reinterpret_cast<const float4*>(p.w + i);
    __syncthreads();
  }
  const float* W = p.w_in_smem ? wsm : p.w;
  const int cgroups = p.CO / 8;
  const int Hs = p.pool ? p.Ho / 2 : p.Ho, Ws = p.pool ? p.Wo / 2 : p.Wo;
  const long long total = (long long)p.n_img * Hs * Ws * cgroups;
  const unsigned int step = p.hb ? p.hb->iterations : 0u, seed = p.hb ? p.hb->seed : 0u;
  const bool do_drop = p.drop_rate > 0.f && p.hb && p.hb->training;
  const float keep_scale = do_drop ? 1.f / (1.f - p.drop_rate) : 1.f;
  const int npos = p.pool ? 4 : 1;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int cg = (int)(r % cgroups); r /= cgroups;
    const int ws = (int)(r % Ws); r /= Ws;
    const int hs = (int)(r % Hs); const int n = (int)(r / Hs);
    const int c0 = cg * 8;
    float acc[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[q][j] = 0.f;
    const float* xn = p.x + (size_t)n * p.Hi * p.Wi * p.CI;
    for (int kh = 0; kh < p.KH; ++kh)
      for (int kw = 0; kw < p.KW; ++kw) {
        const float* wt = W + (size_t)((kh * p.KW + kw) * p.CI) * p.CO + c0;
        for (int ci = 0; ci < p.CI; ++ci) {
          const float4 w0 = *reinterpret_cast<const float4*>(wt + (size_t)ci * p.CO);
          const float4 w1 = *reinterpret_cast<const float4*>(wt + (size_t)ci * p.CO + 4);
          const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (q < npos) {
              const int ho = p.pool ? 2 * hs + (q >> 1) : hs, wo = p.pool ? 2 * ws + (q & 1) : ws;
              const int hi = ho * p.stride + kh - p.pad, wi = wo * p.stride + kw - p.pad;
              float xv = 0.f;
              if (hi >= 0 && hi < p.Hi && wi >= 0 && wi < p.Wi) xv = __ldg(xn + ((size_t)hi * p.Wi + wi) * p.CI + ci);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[q][j] = fmaf(xv, wv[j], acc[q][j]);
            }
          }
        }
      }
    uint32_t o[8];
    const size_t oidx = (((size_t)n * Hs + hs) * Ws + ws) * p.CO + c0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float b = __ldg(p.bias + c0 + j);
      float best = fmaxf(acc[0][j] + b, 0.f); uint32_t pos = 0;
      if (p.pool) {
#pragma unroll
        for (int q = 1; q < 4; ++q) { const float v = fmaxf(acc[q][j] + b, 0.f); if (v > best) { best = v; pos = q; } }
      }
      if (do_drop) {
        const float u = dropout_uniform(seed, step, p.layer_id, (uint32_t)(oidx + j));
        best = u >= p.drop_rate ? best * keep_scale : 0.f;
      }
      o[j] = (__float_as_uint(best) & ~3u) | pos;
    }
    uint4* dst = reinterpret_cast<uint4*>(p.out + oidx);
    dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
    dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
  }
}

// =============================================================================================
// Weight gradient of the first conv layer from the *masked* gradient at activation resolution
// (the arg-max bits in the stored activation say which pre-pool pixel each gradient belongs to).
//   g   : [n][Hs][Ws][CO]  gradient w.r.t. the stored activation, already masked and dropout-scaled
//   act : [n][Hs][Ws][CO]  stored activation (2 LSBs = arg-max position) -- only read when pool = 1
// Thread = (channel co, pixel lane); 9*CI partial sums live in registers, reduced through smem, then red.add.
// =============================================================================================
struct WgradFirstParams {
  const float* x; const float* g; const float* act;
  float* dW; float* db;
  int n_img, Hi, Wi, CI, CO, Hs, Ws, pool, pad, stride;
};

template <int CI>
__global__ void __launch_bounds__(256) conv_wgrad_first_kernel(const WgradFirstParams p) {
  extern __shared__ float red[];           // [groups][9*CI + 1][CO]
  const int CO = p.CO;
  const int groups = blockDim.x / CO;
  const int co = threadIdx.x % CO, grp = threadIdx.x / CO;
  float acc[9 * CI]; float bacc = 0.f;
#pragma unroll
  for (int i = 0; i < 9 * CI; ++i) acc[i] = 0.f;
  const long long npix = (long long)p.n_img * p.Hs * p.Ws;
  if (grp < groups) {
    for (long long pix = (long long)blockIdx.x * groups + grp; pix < npix; pix += (long long)gridDim.x * groups) {
      const float gv = __ldg(p.g + pix * CO + co);
      if (gv == 0.f) continue;
      long long r = pix;
      const int ws = (int)(r % p.Ws); r /= p.Ws;
      const int hs = (int)(r % p.Hs); const int n = (int)(r / p.Hs);
      int ho = hs, wo = ws;
      if (p.pool) {
        const uint32_t pos = __float_as_uint(__ldg(p.act + pix * CO + co)) & 3u;
        ho = 2 * hs + (pos >> 1); wo = 2 * ws + (pos & 1);
      }
      bacc += gv;
      const float* xn = p.x + (size_t)n * p.Hi * p.Wi * CI;
#pragma unroll
      for (int kh = 0; kh < 3; ++kh) {
        const int hi = ho * p.stride + kh - p.pad;
        if (hi < 0 || hi >= p.Hi) continue;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const int wi = wo * p.stride + kw - p.pad;
          if (wi < 0 || wi >= p.Wi) continue;
          const float* xp = xn + ((size_t)hi * p.Wi + wi) * CI;
#pragma unroll
          for (int ci = 0; ci < CI; ++ci) acc[(kh * 3 + kw) * CI + ci] = fmaf(__ldg(xp + ci), gv, acc[(kh * 3 + kw) * CI + ci]);
        }
      }
    }
  }
  const int R = 9 * CI + 1;
  if (grp < groups) {
#pragma unroll
    for (int i = 0; i < 9 * CI; ++i) red[((size_t)grp * R + i) * CO + co] = acc[i];
    red[((size_t)grp * R + 9 * CI) * CO + co] = bacc;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < R * CO; i += blockDim.x) {
    float s = 0.f;
    for (int gidx = 0; gidx < groups; ++gidx) s += red[(size_t)gidx * R * CO + i];
    if (s != 0.f) {
      const int row = i / CO, c = i % CO;
      if (row < 9 * CI) red_add_f32(p.dW + (size_t)row * CO + c, s);
      else red_add_f32(p.db + c, s);
    }
  }
}

// =============================================================================================
// First conv layer, shared-memory tiled (3x3, stride 1, CI <= 4, CO <= 64, 2x2 pool).  One block = one
// (16 x 16)-pixel output tile of one image: the (18 x 18 x CI) input halo and the filter live in shared memory, so
// every global value is read once.  The direct kernel above stays as the general fall-back.
// =============================================================================================
struct ConvFirstParams {
  const float* x; const float* w; const float* bias; float* out;     // out: pooled [n][Hs][Ws][CO]
  int n_img, Hi, Wi, CI, CO, Ho, Wo, pad;
  int tiles_w, tiles_h;
  float drop_rate; unsigned int layer_id; const HyperBlock* hb;
  // backward (wgrad): g = masked gradient at the pooled grid, act = stored activation (arg-max bits)
  const float* g; const float* act; float* dW; float* db;
};

#ifndef CIB_CFF_MINBLOCKS
#define CIB_CFF_MINBLOCKS 6
#endif
template <int CI>
__global__ void __launch_bounds__(128, CIB_CFF_MINBLOCKS) conv_first_fwd_kernel(const ConvFirstParams p) {
  __shared__ float xs[18 * 18 * CI];
  __shared__ float ws[9 * CI * 64];
  __shared__ float bs[64];
  const int CO = p.CO;
  pdl_launch();
  for (int i = threadIdx.x; i < 9 * CI * CO; i += blockDim.x) ws[i] = p.w[i];     // parameters: constant within a step
  for (int i = threadIdx.x; i < CO; i += blockDim.x) bs[i] = p.bias[i];
  pdl_wait();
  const int Hs = p.Ho >> 1, Ws = p.Wo >> 1;
  const unsigned int step = p.hb ? p.hb->iterations : 0u, seed = p.hb ? p.hb->seed : 0u;
  const bool do_drop = p.drop_rate > 0.f && p.hb && p.hb->training;
  const float keep_scale = do_drop ? 1.f / (1.f - p.drop_rate) : 1.f;
  const int ntiles = p.n_img * p.tiles_h * p.tiles_w;
  const int cgroups = CO / 8;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int t = tile;
    const int tw = t % p.tiles_w; t /= p.tiles_w;
    const int th = t % p.tiles_h; const int n = t / p.tiles_h;
    const int h0 = th * 16, w0 = tw * 16;
    __syncthreads();
    const float* xn = p.x + (size_t)n * p.Hi * p.Wi * CI;
    {
      // A halo row is 18 * CI CONTIGUOUS floats in memory: thread r * ROWF + t loads float t of row (pass * RPP + r).  No divisions per
      // element (they were most of this kernel's instructions), one column test per thread and tile.
      constexpr int ROWF = 18 * CI, RPP = 128 / ROWF;           // floats per halo row, rows per pass of the 128 threads
      const int r = threadIdx.x / ROWF, t = threadIdx.x - r * ROWF;
      const int gw = w0 - p.pad + t / CI;
      const bool col_ok = r < RPP && gw >= 0 && gw < p.Wi;
      const float* src = xn + ((long long)(h0 - p.pad) * p.Wi + (w0 - p.pad)) * CI + t;     // only dereferenced where in bounds
#pragma unroll
      for (int hh0 = 0; hh0 < 18; hh0 += RPP) {
        const int hh = hh0 + r, gh = h0 - p.pad + hh;
        if (r < RPP && hh < 18) xs[hh * ROWF + t] = (col_ok && gh >= 0 && gh < p.Hi) ? __ldg(src + (long long)hh * p.Wi * CI) : 0.f;
      }
    }
    __syncthreads();
    // work item = (pooled pixel of the 8x8 pooled tile, group of 8 output channels)
    for (int item = threadIdx.x; item < 64 * cgroups; item += blockDim.x) {
      const int cg = item % cgroups, pp = item / cgroups;
      const int ph = pp >> 3, pw = pp & 7;
      const int c0 = cg * 8;
      float acc[4][8];
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[q][j] = 0.f;
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw)
#pragma unroll
          for (int ci = 0; ci < CI; ++ci) {
            const float4 wa = *reinterpret_cast<const float4*>(&ws[((kh * 3 + kw) * CI + ci) * CO + c0]);
            const float4 wb = *reinterpret_cast<const float4*>(&ws[((kh * 3 + kw) * CI + ci) * CO + c0 + 4]);
            const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float xv = xs[((2 * ph + (q >> 1) + kh) * 18 + 2 * pw + (q & 1) + kw) * CI + ci];
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[q][j] = fmaf(xv, wv[j], acc[q][j]);
            }
          }
      const int gph = (h0 >> 1) + ph, gpw = (w0 >> 1) + pw;
      if (gph < Hs && gpw < Ws) {
        const size_t oidx = (((size_t)n * Hs + gph) * Ws + gpw) * CO + c0;
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float b = bs[c0 + j];
          float best = fmaxf(acc[0][j] + b, 0.f); uint32_t pos = 0;
#pragma unroll
          for (int q = 1; q < 4; ++q) { const float v = fmaxf(acc[q][j] + b, 0.f); if (v > best) { best = v; pos = q; } }
          if (do_drop) best = dropout_uniform(seed, step, p.layer_id, (uint32_t)(oidx + j)) >= p.drop_rate ? best * keep_scale : 0.f;
          o[j] = (__float_as_uint(best) & ~3u) | pos;
        }
        uint4* dst = reinterpret_cast<uint4*>(p.out + oidx);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
      }
    }
  }
}

// wgrad of the first layer, same tiling.  Thread = (output channel co, pixel lane): per pooled pixel it reads the
// gradient and the arg-max position once and does all 9*CI FMAs with compile-time offsets into register accumulators
// (2.3 instructions per FMA instead of 12 for a thread-per-weight mapping).  Sums persist across the tiles a thread sees.
//
// One 1024-thread block per SM made of four independent 256-thread groups, each walking its own tiles with its own
// named barrier: the four groups hide each other's global-load latency like four resident blocks would, but the block
// ends with ONE set of red.adds.  (All blocks add to the same 9*CI*CO+CO addresses, which sit in a handful of L2 lines
// and serialise there: 592 small blocks cost ~20 us of tail, one block per SM with a single group stalls on every tile.)
#define CFW_GROUPS 4
__device__ __forceinline__ void group_bar(int g) { asm volatile("bar.sync %0, 256;" ::"r"(g + 1) : "memory"); }

template <int CI>
__global__ void __launch_bounds__(256 * CFW_GROUPS) conv_first_wgrad_kernel(const ConvFirstParams p) {
  extern __shared__ __align__(16) float cfw_smem[];
  const int CO = p.CO;
  const int grp = threadIdx.x >> 8, tid = threadIdx.x & 255;
  float* xs = cfw_smem + grp * (18 * 18 * CI);                                   // [CFW_GROUPS][18*18*CI]
  float* gs = cfw_smem + CFW_GROUPS * (18 * 18 * CI) + grp * (64 * CO);          // [CFW_GROUPS][64 pooled pixels][CO]
  const int lanes = 256 / CO;                        // pixel lanes per channel
  const int co = tid % CO, pl = tid / CO;
  const int Hs = p.Ho >> 1, Ws = p.Wo >> 1;
  float acc[9 * CI]; float bacc = 0.f;
#pragma unroll
  for (int i = 0; i < 9 * CI; ++i) acc[i] = 0.f;
  constexpr int R = 9 * CI + 1;
  pdl_launch();
  pdl_wait();
  const int ntiles = p.n_img * p.tiles_h * p.tiles_w;
  for (int tile = blockIdx.x * CFW_GROUPS + grp; tile < ntiles; tile += gridDim.x * CFW_GROUPS) {
    int t = tile;
    const int tw = t % p.tiles_w; t /= p.tiles_w;
    const int th = t % p.tiles_h; const int n = t / p.tiles_h;
    const int h0 = th * 16, w0 = tw * 16;
    group_bar(grp);
    const float* xn = p.x + (size_t)n * p.Hi * p.Wi * CI;
    {
      // halo rows are contiguous in memory (see the forward kernel): thread r * ROWF + t loads float t of row (pass * RPP + r)
      constexpr int ROWF = 18 * CI, RPP = 256 / ROWF;
      const int r = tid / ROWF, t = tid - r * ROWF;
      const int gw = w0 - p.pad + t / CI;
      const bool col_ok = r < RPP && gw >= 0 && gw < p.Wi;
      const float* src = xn + ((long long)(h0 - p.pad) * p.Wi + (w0 - p.pad)) * CI + t;
#pragma unroll
      for (int hh0 = 0; hh0 < 18; hh0 += RPP) {
        const int hh = hh0 + r, gh = h0 - p.pad + hh;
        if (r < RPP && hh < 18) xs[hh * ROWF + t] = (col_ok && gh >= 0 && gh < p.Hi) ? __ldg(src + (long long)hh * p.Wi * CI) : 0.f;
      }
    }
    for (int i = tid; i < 64 * CO; i += 256) {                           // coalesced: 8 pooled pixels x CO are contiguous
      const int c = i % CO, pp = i / CO;
      const int gph = (h0 >> 1) + (pp >> 3), gpw = (w0 >> 1) + (pp & 7);
      uint32_t bits = 0u;
      if (gph < Hs && gpw < Ws) {
        const size_t idx = (((size_t)n * Hs + gph) * Ws + gpw) * CO + c;
        bits = (__float_as_uint(__ldg(p.g + idx)) & ~3u) | (__float_as_uint(__ldg(p.act + idx)) & 3u);
      }
      gs[i] = __uint_as_float(bits);
    }
    group_bar(grp);
    if (pl < lanes) {
      for (int pp = pl; pp < 64; pp += lanes) {
        const uint32_t bits = __float_as_uint(gs[pp * CO + co]);
        const float gv = __uint_as_float(bits & ~3u);
        if (gv == 0.f) continue;
        const int pos = (int)(bits & 3u);
        bacc += gv;
        const float* xb = &xs[((2 * (pp >> 3) + (pos >> 1)) * 18 + 2 * (pp & 7) + (pos & 1)) * CI];
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw)
#pragma unroll
            for (int ci = 0; ci < CI; ++ci)
              acc[(kh * 3 + kw) * CI + ci] = fmaf(xb[(kh * 18 + kw) * CI + ci], gv, acc[(kh * 3 + kw) * CI + ci]);
      }
    }
  }
  __syncthreads();                                   // every group is done with its tiles: the tile buffers can be reused
  // Block reduction WITHOUT shared-memory atomics (28 rows x 32 warps of same-address atomicAdd serialised for ~10 us): lanes of a
  // warp that hold the same channel (CO < 32) are folded with shuffles, every warp writes its R x CO partials to its own slab of
  // shared memory (the tile buffers are free now), and R x CO threads add the slabs up and issue one red.add each.
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = 8 * CFW_GROUPS;
  float* slab = cfw_smem;                            // [NW][R * CO]  (host sizes the dynamic shared memory for it)
#pragma unroll
  for (int i = 0; i < R; ++i) {
    float v = (pl < lanes) ? (i < 9 * CI ? acc[i < 9 * CI ? i : 0] : bacc) : 0.f;
    for (int off = 16; off >= CO; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (CO >= 32 || lane < CO) {
      slab[(size_t)warp * (R * CO) + i * CO + co] = v;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < R * CO; i += blockDim.x) {
    float sum = 0.f;
    if (CO >= 32) {
      // with CO >= 32 a warp holds only part of the channels: slabs of warps that do not own channel (i % CO) were never written
      const int c = i % CO;
      for (int w = 0; w < NW; ++w) if (((w * 32) % CO) <= c && c < ((w * 32) % CO) + 32) sum += slab[(size_t)w * (R * CO) + i];
    } else {
#pragma unroll 8
      for (int w = 0; w < NW; ++w) sum += slab[(size_t)w * (R * CO) + i];
    }
    if (sum != 0.f) {
      if (i < 9 * CI * CO) red_add_f32(p.dW + i, sum);
      else red_add_f32(p.db + (i - 9 * CI * CO), sum);
    }
  }
}

// =============================================================================================
// Shape-agnostic fall-backs (any kernel size / stride / channel count).  Slow; used for the strided
// legacy model of Train_rpv.ipynb and as an on-GPU cross-check of the tcgen05 kernels.
// =============================================================================================
struct ConvGenericBwdParams {
  const float* x;        // layer input  [n][Hi][Wi][CI]
  const float* dz;       // gradient w.r.t. the conv output (pre-pool resolution) [n][Ho][Wo][CO]
  const float* w;        // [KH*KW][CI][CO]
  float* dW; float* db;  // wgrad outputs
  // dgrad outputs / epilogue (same conventions as conv_tc.cuh)
  const float* act_prev; float* dx_out; int epi; int Hf, Wf; float scale;
  int n_img, Hi, Wi, CI, CO, Ho, Wo, KH, KW, stride, pad;
};

__global__ void __launch_bounds__(256) conv_dgrad_generic_kernel(const ConvGenericBwdParams p) {
  const long long total = (long long)p.n_img * p.Hi * p.Wi * p.CI;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int ci = (int)(r % p.CI); r /= p.CI;
    const int wi = (int)(r % p.Wi); r /= p.Wi;
    const int hi = (int)(r % p.Hi); const int n = (int)(r / p.Hi);
    float acc = 0.f;
    for (int kh = 0; kh < p.KH; ++kh) {
      const int hn = hi + p.pad - kh; if (hn < 0 || hn % p.stride) continue;
      const int ho = hn / p.stride; if (ho >= p.Ho) continue;
      for (int kw = 0; kw < p.KW; ++kw) {
        const int wn = wi + p.pad - kw; if (wn < 0 || wn % p.stride) continue;
        const int wo = wn / p.stride; if (wo >= p.Wo) continue;
        const float* dzp = p.dz + (((size_t)n * p.Ho + ho) * p.Wo + wo) * p.CO;
        const float* wp = p.w + ((size_t)(kh * p.KW + kw) * p.CI + ci) * p.CO;
        for (int co = 0; co < p.CO; ++co) acc = fmaf(__ldg(dzp + co), __ldg(wp + co), acc);
      }
    }
    const uint32_t ab = __float_as_uint(__ldg(p.act_prev + idx));
    const float g = (ab & ~3u) ? acc * p.scale : 0.f;
    if (p.epi == CONV_EPI_DGRAD_MASK) {
      p.dx_out[idx] = g;
    } else {
      const int pos = ab & 3;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int hh = 2 * hi + (q >> 1), ww = 2 * wi + (q & 1);
        p.dx_out[(((size_t)n * p.Hf + hh) * p.Wf + ww) * p.CI + ci] = (q == pos) ? g : 0.f;
      }
    }
  }
}

// one block per (tap, ci) row; threads over co x pixel lanes
__global__ void __launch_bounds__(256) conv_wgrad_generic_kernel(const ConvGenericBwdParams p) {
  __shared__ float red[256];
  const int row = blockIdx.x;                      // (kh*KW + kw)*CI + ci, or KH*KW*CI for the bias row
  const int nrows = p.KH * p.KW * p.CI;
  const bool bias_row = row == nrows;
  const int ci = row % p.CI; const int tap = row / p.CI;
  const int kh = tap / p.KW, kw = tap % p.KW;
  const long long npix = (long long)p.n_img * p.Ho * p.Wo;
  const int slice = blockIdx.y, nslices = gridDim.y;
  for (int co0 = 0; co0 < p.CO; co0 += 32) {
    const int co = co0 + (threadIdx.x & 31);
    const int lanep = threadIdx.x >> 5;            // 8 pixel lanes
    float acc = 0.f;
    if (co < p.CO) {
      for (long long pix = (long long)slice * 8 + lanep; pix < npix; pix += (long long)nslices * 8) {
        long long r = pix;
        const int wo = (int)(r % p.Wo); r /= p.Wo;
        const int ho = (int)(r % p.Ho); const int n = (int)(r / p.Ho);
        const float dzv = __ldg(p.dz + pix * p.CO + co);
        if (bias_row) { acc += dzv; continue; }
        const int hi = ho * p.stride + kh - p.pad, wi = wo * p.stride + kw - p.pad;
        if (hi < 0 || hi >= p.Hi || wi < 0 || wi >= p.Wi) continue;
        acc = fmaf(__ldg(p.x + (((size_t)n * p.Hi + hi) * p.Wi + wi) * p.CI + ci), dzv, acc);
      }
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
      float s = 0.f;
      for (int l = 0; l < 8; ++l) s += red[l * 32 + threadIdx.x];
      if (co < p.CO && s != 0.f) {
        if (bias_row) red_add_f32(p.db + co, s);
        else red_add_f32(p.dW + (size_t)row * p.CO + co, s);
      }
    }
    __syncthreads();
  }
}

// Expand a masked gradient at pooled resolution to pre-pool resolution (generic path only).
__global__ void unpool_generic_kernel(const float* g, const float* act, float* dz, int n_img, int Hs, int Ws, int C, int Hf, int Wf) {
  const long long total = (long long)n_img * Hs * Ws * C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int c = (int)(r % C); r /= C;
    const int ws = (int)(r % Ws); r /= Ws;
    const int hs = (int)(r % Hs); const int n = (int)(r / Hs);
    const int pos = __float_as_uint(act[idx]) & 3;
    const float gv = g[idx];
#pragma unroll
    for (int q = 0; q < 4; ++q)
      dz[(((size_t)n * Hf + 2 * hs + (q >> 1)) * Wf + 2 * ws + (q & 1)) * C + c] = (q == pos) ? gv : 0.f;
  }
}

struct DenseGenericParams {
  const float* a;      // [M][K]
  const float* w;      // [K][N]
  const float* dz;     // [M][N]
  float* z;            // fwd out [M][N] (accumulated with +=, like the split-K tcgen05 path)
  float* dW;           // [K][N]
  // dgrad epilogue
  const float* act_prev; float* dx_out; int epi; int Hp, Wp, Cp, Hf, Wf; float scale;
  int M, N, K;
};

__global__ void __launch_bounds__(256) dense_fwd_generic_kernel(const DenseGenericParams p) {
  const long long total = (long long)p.M * p.N;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % p.N), m = (int)(idx / p.N);
    float acc = 0.f;
    const float* ap = p.a + (size_t)m * p.K;
    for (int k = 0; k < p.K; ++k) acc = fmaf(__ldg(ap + k), __ldg(p.w + (size_t)k * p.N + n), acc);
    p.z[idx] += acc;
  }
}
__global__ void __launch_bounds__(256) dense_wgrad_generic_kernel(const DenseGenericParams p) {
  const long long total = (long long)p.K * p.N;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % p.N), k = (int)(idx / p.N);
    float acc = 0.f;
    for (int m = 0; m < p.M; ++m) acc = fmaf(__ldg(p.a + (size_t)m * p.K + k), __ldg(p.dz + (size_t)m * p.N + n), acc);
    p.dW[idx] = acc;
  }
}
// dx[m][k] = sum_n dz[m][n] w[k][n], then mask (and optionally un-pool) exactly like the tcgen05 epilogues
__global__ void __launch_bounds__(256) dense_dgrad_generic_kernel(const DenseGenericParams p) {
  const long long total = (long long)p.M * p.K;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(idx % p.K), m = (int)(idx / p.K);
    float acc = 0.f;
    const float* dzp = p.dz + (size_t)m * p.N; const float* wp = p.w + (size_t)k * p.N;
    for (int n = 0; n < p.N; ++n) acc = fmaf(__ldg(dzp + n), __ldg(wp + n), acc);
    const uint32_t ab = __float_as_uint(__ldg(p.act_prev + idx));
    const float g = (ab & ~3u) ? acc * p.scale : 0.f;
    if (p.epi == CONV_EPI_DGRAD_MASK) { p.dx_out[idx] = g; }
    else {
      const int c = k % p.Cp; const int pp = k / p.Cp; const int pw = pp % p.Wp, ph = pp / p.Wp;
      const int pos = ab & 3;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        p.dx_out[(((size_t)m * p.Hf + 2 * ph + (q >> 1)) * p.Wf + 2 * pw + (q & 1)) * p.Cp + c] = (q == pos) ? g : 0.f;
    }
  }
}

// =============================================================================================
// Hidden dense layer activation: a = dropout(relu(z + b)), and z is cleared for the next split-K accumulation.
// =============================================================================================
__global__ void bias_relu_dropout_kernel(float* z, const float* bias, float* a, int M, int N, float drop_rate,
                                         unsigned int layer_id, const HyperBlock* hb) {
  pdl_launch(); pdl_wait();
  const unsigned int step = hb ? hb->iterations : 0u, seed = hb ? hb->seed : 0u;
  const bool do_drop = drop_rate > 0.f && hb && hb->training;
  const float keep_scale = do_drop ? 1.f / (1.f - drop_rate) : 1.f;
  const int total = M * N;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    float v = fmaxf(z[idx] + __ldg(bias + idx % N), 0.f);
    z[idx] = 0.f;
    if (do_drop) v = dropout_uniform(seed, step, layer_id, (uint32_t)idx) >= drop_rate ? v * keep_scale : 0.f;
    a[idx] = __uint_as_float(__float_as_uint(v) & ~3u);
  }
}

// =============================================================================================
// Loss head: output Dense (+bias) + sigmoid/softmax + Keras cross-entropy + accuracy, and in training mode the
// whole backward of the head: dW_out, db_out, and the masked gradient of the last hidden activation together
// with that layer's bias gradient.  One CTA (the head is 128 x 128 x {1,10} MACs); reductions stay in smem
// so results are deterministic.  Reference: Dense(1,'sigmoid')+binary_crossentropy rpv.py:59,70;
// Dense(10,'softmax')+categorical_crossentropy mnist.py:56-57; metric 'accuracy' mnist.py:58.
// =============================================================================================
struct HeadParams {
  const float* a;        // [B][F] last hidden activation (post ReLU/dropout)
  const float* w;        // [F][O]
  const float* bias;     // [O]
  const float* y;        // [B][O] targets
  const float* sw;       // [B] sample weights (0 for padding rows), may be null
  float* probs;          // [B][O] (may be null)
  float* dW; float* db;  // [F][O], [O]
  float* dz_hidden;      // [B][F] masked gradient w.r.t. the last hidden pre-activation
  float* db_hidden;      // [F]
  HyperBlock* hb;
  int B, F, O, kind;     // kind 0 = sigmoid + binary CE, 1 = softmax + categorical CE
  int train;
  int a_in_smem;       // 1 = the [B][F] activation tile is staged in shared memory (fits), 0 = read from global
  int advance_cursor;  // 1 = this step consumed B rows of the device-resident permutation
  float hidden_scale;    // 1/(1-rate) of the dropout after the last hidden layer
  // Fused activation of the last hidden Dense layer (multi-CTA head only): when z is set, `a` has not been computed yet -- the head
  // reads the split-K accumulator z [B][F], applies bias + ReLU + dropout, stores the activation to act_out (the backward mask),
  // puts z back to zero for the next step and goes on with the values in shared memory.  One launch fewer per step.
  float* z; const float* bias_hidden; float* act_out; float drop_hidden; unsigned int layer_hidden;
  // Host-fed training: the step's {loss, accuracy} also go straight to a ring in pinned, device-mapped HOST memory (slot = optimizer
  // step & ring_mask) -- the device-to-host read of every step's result without a copy node between two step graphs.
  float* loss_ring; unsigned int ring_mask;
};

#define CIB_HEAD_MAXO 16
// OM = compile-time bound of the output-unit loops (1 for the sigmoid head, 16 for softmax heads): with a runtime bound
// every loop body was emitted 16 times under predicates (162 k warp instructions for a 128 x 128 x 1 head, measured).
template <int OM>
__global__ void __launch_bounds__(1024, 1) head_kernel(const HeadParams p) {
  extern __shared__ float hs[];                     // dl[B][O] | red[128] | a[B][F] (optional)
  pdl_launch(); pdl_wait();
  float* dl = hs;                                   // dLoss/dlogit (already divided by the batch weight sum)
  float* red = hs + ((p.B * p.O + 3) & ~3);         // block reduction scratch (16-byte aligned)
  float* a_s = red + 128;
  if (p.a_in_smem) {
    // every later phase walks the batch dimension serially per thread: from global memory that is 128 dependent
    // ~0.5 us loads (measured 51 us for the whole kernel); staged in shared memory it is ~30 cycles each
    const int n4 = p.B * p.F / 4;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) reinterpret_cast<float4*>(a_s)[i] = reinterpret_cast<const float4*>(p.a)[i];
    __syncthreads();
  }
  const float* A = p.a_in_smem ? a_s : p.a;
  __shared__ float s_wsum;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int B = p.B, F = p.F, O = p.O;
  const float eps = 1e-7f;

  // total sample weight (denominator of the batch mean)
  float wl = 0.f;
  for (int m = threadIdx.x; m < B; m += blockDim.x) wl += p.sw ? p.sw[m] : 1.f;
  for (int o = 16; o; o >>= 1) wl += __shfl_xor_sync(0xffffffffu, wl, o);
  if (lane == 0) red[warp] = wl;
  __syncthreads();
  if (threadIdx.x == 0) { float s = 0.f; for (int i = 0; i < nwarps; ++i) s += red[i]; s_wsum = s; }
  __syncthreads();
  const float wsum = fmaxf(s_wsum, 1e-20f);

  float loss_acc = 0.f, corr_acc = 0.f;
  for (int m = warp; m < B; m += nwarps) {
    float logit[OM];
#pragma unroll
    for (int o = 0; o < OM; ++o) logit[o] = 0.f;
    for (int f = lane; f < F; f += 32) {
      const float av = A[(size_t)m * F + f];
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) logit[o] = fmaf(av, __ldg(p.w + (size_t)f * O + o), logit[o]);
    }
#pragma unroll
    for (int o = 0; o < OM; ++o)
      if (o < O) { for (int s = 16; s; s >>= 1) logit[o] += __shfl_xor_sync(0xffffffffu, logit[o], s); logit[o] += __ldg(p.bias + o); }
    const float swm = p.sw ? p.sw[m] : 1.f;
    float loss = 0.f, correct = 0.f;
    float prob[OM], grad[OM];
    if (p.kind == 0) {
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) {
        const float pr = 1.f / (1.f + __expf(-logit[o]));
        const float yv = p.y[(size_t)m * O + o];
        const float pc = fminf(fmaxf(pr, eps), 1.f - eps);
        loss += -(yv * __logf(pc) + (1.f - yv) * __logf(1.f - pc));
        correct += ((pr > 0.5f ? 1.f : 0.f) == yv) ? 1.f : 0.f;
        prob[o] = pr;
        grad[o] = (pc == pr) ? (pr - yv) / O : 0.f;
      }
      loss /= O; correct /= O;
    } else {
      float mx = -INFINITY;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) mx = fmaxf(mx, logit[o]);
      float sum = 0.f;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) { prob[o] = __expf(logit[o] - mx); sum += prob[o]; }
      int am = 0, ay = 0; float bp = -1.f, by = -1.f; bool clipped = false;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) {
        prob[o] /= sum;
        const float yv = p.y[(size_t)m * O + o];
        const float pc = fminf(fmaxf(prob[o], eps), 1.f - eps);
        loss += -yv * __logf(pc);
        if (yv > 0.f && pc != prob[o]) clipped = true;
        if (prob[o] > bp) { bp = prob[o]; am = o; }
        if (yv > by) { by = yv; ay = o; }
      }
      correct = (am == ay) ? 1.f : 0.f;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) grad[o] = clipped ? 0.f : prob[o] - p.y[(size_t)m * O + o];
    }
    if (lane == 0) {
      loss_acc += swm * loss; corr_acc += swm * correct;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) {
        dl[m * O + o] = grad[o] * swm / wsum;
        if (p.probs) p.probs[(size_t)m * O + o] = prob[o];
      }
    }
  }
  // block-reduce loss / correct
  if (lane == 0) { red[warp] = loss_acc; red[32 + warp] = corr_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f;
    for (int i = 0; i < nwarps; ++i) { l += red[i]; c += red[32 + i]; }
    p.hb->loss_sum += l; p.hb->correct_sum += c; p.hb->sample_sum += s_wsum;
    p.hb->step_loss = l / wsum; p.hb->step_acc = c / wsum;
    if (p.loss_ring && p.train) {
      float* slot = p.loss_ring + 2 * (size_t)(p.hb->iterations & p.ring_mask);
      slot[0] = l / wsum; slot[1] = c / wsum;
    }
    if (p.advance_cursor) advance_batch_cursor(p.hb, B);
  }
  if (!p.train) return;
  __syncthreads();

  // ---- backward of the head, all 1024 threads: thread = (feature f, batch group mg).  Each walks B/G samples and keeps
  //      partial dW_out[f][:], db_hidden[f]; a shared-memory reduction over the G groups finishes.  (A thread-per-output
  //      mapping left 4 of 32 warps running 128-iteration dependent loops: 94 k cycles, measured.)
  const int G = blockDim.x / F;                      // batch groups (>= 1 since F <= 256... checked by the launcher)
  float* part = A == a_s ? a_s + (size_t)B * F : red + 128;   // [G][F][O + 1] partial sums
  const int f = threadIdx.x % F, mg = threadIdx.x / F;
  if (mg < G) {
    float wrow[OM], dwp[OM];
#pragma unroll
    for (int o = 0; o < OM; ++o) { wrow[o] = o < O ? __ldg(p.w + (size_t)f * O + o) : 0.f; dwp[o] = 0.f; }
    float bsum = 0.f;
    for (int m = mg; m < B; m += G) {
      const float av = A[(size_t)m * F + f];
      float g = 0.f;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) { const float d = dl[m * O + o]; dwp[o] = fmaf(av, d, dwp[o]); g = fmaf(d, wrow[o], g); }
      g = (__float_as_uint(av) & ~3u) ? g * p.hidden_scale : 0.f;
      p.dz_hidden[(size_t)m * F + f] = g;
      bsum += g;
    }
#pragma unroll
    for (int o = 0; o < OM; ++o) if (o < O) part[((size_t)mg * F + f) * (O + 1) + o] = dwp[o];
    part[((size_t)mg * F + f) * (O + 1) + O] = bsum;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < F * (O + 1); i += blockDim.x) {
    const int ff = i / (O + 1), o = i % (O + 1);
    float sum = 0.f;
    for (int g2 = 0; g2 < G; ++g2) sum += part[((size_t)g2 * F + ff) * (O + 1) + o];
    if (o < O) p.dW[ff * O + o] += sum; else p.db_hidden[ff] += sum;
  }
  for (int o = threadIdx.x; o < O; o += blockDim.x) {
    float sum = 0.f;
    for (int m = 0; m < B; ++m) sum += dl[m * O + o];
    p.db[o] += sum;
  }
}

// Multi-CTA version of the loss head.  The single-CTA kernel above is latency-bound (19 us for the 128 x 128 x 1 sigmoid head,
// 60 us for a 128 x 128 x 10 softmax head, on the critical path of every step while 147 SMs idle): here CTA c takes a slice of
// the batch, does forward + loss + its rows of dz_hidden, and adds its partial dW / db / db_hidden / metric sums with red.add
// (the gradients are accumulation buffers anyway).  The last CTA to finish (ticket in `scratch`) publishes step_loss / step_acc
// and advances the batch cursor; it sums the per-CTA loss / accuracy partials in CTA order, so the reported metrics are
// deterministic.  scratch = {loss partial[64], correct partial[64], ticket}, ticket zero between launches.
template <int OM>
__global__ void __launch_bounds__(256) head_mc_kernel(const HeadParams p, float* scratch) {
  extern __shared__ float hs[];
  pdl_launch(); pdl_wait();
  const int B = p.B, F = p.F, O = p.O;
  const int S = (B + gridDim.x - 1) / gridDim.x;          // samples per CTA
  const int m0 = blockIdx.x * S, m1 = min(B, m0 + S), ns = max(0, m1 - m0);
  float* a_s = hs;                                        // [S][F]
  float* w_s = a_s + (size_t)S * F;                       // [F][O]
  float* dl = w_s + (size_t)F * O;                        // [S][O]
  float* red = dl + ((S * O + 3) & ~3);                   // [64]
  __shared__ float s_wsum;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const float eps = 1e-7f;
  if (p.z) {
    const unsigned int step = p.hb->iterations, seed = p.hb->seed;
    const bool do_drop = p.drop_hidden > 0.f && p.hb->training;
    const float keep_scale = do_drop ? 1.f / (1.f - p.drop_hidden) : 1.f;
    for (int i = threadIdx.x; i < ns * F; i += blockDim.x) {
      const size_t gi = (size_t)m0 * F + i;
      float v = fmaxf(__ldcg(p.z + gi) + __ldg(p.bias_hidden + i % F), 0.f);
      p.z[gi] = 0.f;
      if (do_drop) v = dropout_uniform(seed, step, p.layer_hidden, (uint32_t)gi) >= p.drop_hidden ? v * keep_scale : 0.f;
      v = __uint_as_float(__float_as_uint(v) & ~3u);
      p.act_out[gi] = v;
      a_s[i] = v;
    }
  } else {
    for (int i = threadIdx.x; i < ns * F; i += blockDim.x) a_s[i] = p.a[(size_t)m0 * F + i];
  }
  for (int i = threadIdx.x; i < F * O; i += blockDim.x) w_s[i] = p.w[i];
  float wl = 0.f;
  for (int m = threadIdx.x; m < B; m += blockDim.x) wl += p.sw ? p.sw[m] : 1.f;
  for (int o = 16; o; o >>= 1) wl += __shfl_xor_sync(0xffffffffu, wl, o);
  if (lane == 0) red[warp] = wl;
  __syncthreads();
  if (threadIdx.x == 0) { float s = 0.f; for (int i = 0; i < nwarps; ++i) s += red[i]; s_wsum = s; }
  __syncthreads();
  const float wsum = fmaxf(s_wsum, 1e-20f);

  float loss_acc = 0.f, corr_acc = 0.f;
  for (int ml = warp; ml < ns; ml += nwarps) {
    const int m = m0 + ml;
    float logit[OM];
#pragma unroll
    for (int o = 0; o < OM; ++o) logit[o] = 0.f;
    for (int f = lane; f < F; f += 32) {
      const float av = a_s[ml * F + f];
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) logit[o] = fmaf(av, w_s[f * O + o], logit[o]);
    }
#pragma unroll
    for (int o = 0; o < OM; ++o)
      if (o < O) { for (int s = 16; s; s >>= 1) logit[o] += __shfl_xor_sync(0xffffffffu, logit[o], s); logit[o] += __ldg(p.bias + o); }
    const float swm = p.sw ? p.sw[m] : 1.f;
    float loss = 0.f, correct = 0.f;
    float prob[OM], grad[OM];
    if (p.kind == 0) {
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) {
        const float pr = 1.f / (1.f + __expf(-logit[o]));
        const float yv = p.y[(size_t)m * O + o];
        const float pc = fminf(fmaxf(pr, eps), 1.f - eps);
        loss += -(yv * __logf(pc) + (1.f - yv) * __logf(1.f - pc));
        correct += ((pr > 0.5f ? 1.f : 0.f) == yv) ? 1.f : 0.f;
        prob[o] = pr;
        grad[o] = (pc == pr) ? (pr - yv) / O : 0.f;
      }
      loss /= O; correct /= O;
    } else {
      float mx = -INFINITY;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) mx = fmaxf(mx, logit[o]);
      float sum = 0.f;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) { prob[o] = __expf(logit[o] - mx); sum += prob[o]; }
      int am = 0, ay = 0; float bp = -1.f, by = -1.f; bool clipped = false;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) {
        prob[o] /= sum;
        const float yv = p.y[(size_t)m * O + o];
        const float pc = fminf(fmaxf(prob[o], eps), 1.f - eps);
        loss += -yv * __logf(pc);
        if (yv > 0.f && pc != prob[o]) clipped = true;
        if (prob[o] > bp) { bp = prob[o]; am = o; }
        if (yv > by) { by = yv; ay = o; }
      }
      correct = (am == ay) ? 1.f : 0.f;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) grad[o] = clipped ? 0.f : prob[o] - p.y[(size_t)m * O + o];
    }
    if (lane == 0) {
      loss_acc += swm * loss; corr_acc += swm * correct;
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O) {
        dl[ml * O + o] = grad[o] * swm / wsum;
        if (p.probs) p.probs[(size_t)m * O + o] = prob[o];
      }
    }
  }
  if (lane == 0) { red[warp] = loss_acc; red[32 + warp] = corr_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f;
    for (int i = 0; i < nwarps; ++i) { l += red[i]; c += red[32 + i]; }
    scratch[blockIdx.x] = l; scratch[64 + blockIdx.x] = c;
  }
  if (p.train) {
    // backward for this CTA's samples: thread = feature f (strided); no cross-thread reduction needed
    for (int f = threadIdx.x; f < F; f += blockDim.x) {
      float wrow[OM], dwp[OM];
#pragma unroll
      for (int o = 0; o < OM; ++o) { wrow[o] = o < O ? w_s[f * O + o] : 0.f; dwp[o] = 0.f; }
      float bsum = 0.f;
      for (int ml = 0; ml < ns; ++ml) {
        const float av = a_s[ml * F + f];
        float g = 0.f;
#pragma unroll
        for (int o = 0; o < OM; ++o) if (o < O) { const float d = dl[ml * O + o]; dwp[o] = fmaf(av, d, dwp[o]); g = fmaf(d, wrow[o], g); }
        g = (__float_as_uint(av) & ~3u) ? g * p.hidden_scale : 0.f;
        p.dz_hidden[(size_t)(m0 + ml) * F + f] = g;
        bsum += g;
      }
#pragma unroll
      for (int o = 0; o < OM; ++o) if (o < O && dwp[o] != 0.f) red_add_f32(p.dW + f * O + o, dwp[o]);
      if (bsum != 0.f) red_add_f32(p.db_hidden + f, bsum);
    }
    for (int o = threadIdx.x; o < O; o += blockDim.x) {
      float sum = 0.f;
      for (int ml = 0; ml < ns; ++ml) sum += dl[ml * O + o];
      if (sum != 0.f) red_add_f32(p.db + o, sum);
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 128);
    const unsigned int tk = atomicAdd(ticket, 1u);
    if (tk == gridDim.x - 1) {
      __threadfence();
      float l = 0.f, c = 0.f;
      for (unsigned int i = 0; i < gridDim.x; ++i) { l += __ldcg(scratch + i); c += __ldcg(scratch + 64 + i); }
      *ticket = 0u;
      p.hb->loss_sum += l; p.hb->correct_sum += c; p.hb->sample_sum += s_wsum;
      p.hb->step_loss = l / wsum; p.hb->step_acc = c / wsum;
      if (p.loss_ring && p.train) {
        float* slot = p.loss_ring + 2 * (size_t)(p.hb->iterations & p.ring_mask);
        slot[0] = l / wsum; slot[1] = c / wsum;
      }
      if (p.advance_cursor) advance_batch_cursor(p.hb, B);
      __threadfence();
    }
  }
}

// column sums of a [M][N] matrix added into out[N] (bias gradient of non-final hidden dense layers)
__global__ void colsum_add_kernel(const float* g, float* out, int M, int N) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.f;
  for (int m = 0; m < M; ++m) s += g[(size_t)m * N + n];
  out[n] += s;
}

// =============================================================================================
// Multi-tensor optimizer over the flat parameter buffer (Keras 2.2 update rules, eps outside the sqrt).
// kind: 0 SGD(momentum), 1 Adam, 2 Nadam, 3 Adadelta.  Gradients are cleared after use so the wgrad kernels
// can keep accumulating with red.add; the last block to finish advances the step counter.
// Reference: optimizer-by-name compile at mnist.py:57, rpv.py:62.
// =============================================================================================
struct OptCoef { float lr_t, c1, c2, c3, c4; };

__device__ __forceinline__ OptCoef opt_coefficients(const HyperBlock* hb, int kind) {
  OptCoef c{};
  const float t = (float)(hb->iterations + 1u);
  float lr = hb->lr;
  if (hb->decay > 0.f) lr *= 1.f / (1.f + hb->decay * (float)hb->iterations);
  c.lr_t = lr;
  if (kind == 1) {
    c.lr_t = lr * sqrtf(1.f - powf(hb->beta2, t)) / (1.f - powf(hb->beta1, t));
  } else if (kind == 2) {
    const float mu_t = hb->beta1 * (1.f - 0.5f * powf(0.96f, t * hb->schedule_decay));
    const float mu_t1 = hb->beta1 * (1.f - 0.5f * powf(0.96f, (t + 1.f) * hb->schedule_decay));
    const float sched = hb->m_schedule[hb->iterations & 1u] * mu_t;
    const float sched_next = sched * mu_t1;
    c.c1 = 1.f / (1.f - sched);             // g' = g * c1
    c.c2 = mu_t1 / (1.f - sched_next);      // m' contribution
    c.c3 = 1.f - mu_t;
    c.c4 = 1.f / (1.f - powf(hb->beta2, t));
  }
  return c;
}

__device__ __forceinline__ void opt_update(int kind, const HyperBlock* hb, const OptCoef& c, float g, float& pv, float& s0, float& s1) {
  if (kind == 0) {
    s0 = hb->momentum * s0 - c.lr_t * g; pv += s0;
  } else if (kind == 1) {
    s0 = hb->beta1 * s0 + (1.f - hb->beta1) * g;
    s1 = hb->beta2 * s1 + (1.f - hb->beta2) * g * g;
    pv -= c.lr_t * s0 / (sqrtf(s1) + hb->eps);
  } else if (kind == 2) {
    const float gp = g * c.c1;
    s0 = hb->beta1 * s0 + (1.f - hb->beta1) * g;
    s1 = hb->beta2 * s1 + (1.f - hb->beta2) * g * g;
    const float mbar = c.c3 * gp + c.c2 * s0;
    pv -= c.lr_t * mbar / (sqrtf(s1 * c.c4) + hb->eps);
  } else {
    s0 = hb->rho * s0 + (1.f - hb->rho) * g * g;
    const float u = g * sqrtf(s1 + hb->eps) / sqrtf(s0 + hb->eps);
    pv -= c.lr_t * u;
    s1 = hb->rho * s1 + (1.f - hb->rho) * u * u;
  }
}

__device__ __forceinline__ void opt_advance_state(HyperBlock* hb, int kind) {
  if (kind == 2) {
    const float t = (float)(hb->iterations + 1u);
    const float mu_t = hb->beta1 * (1.f - 0.5f * powf(0.96f, t * hb->schedule_decay));
    hb->m_schedule[(hb->iterations + 1u) & 1u] = hb->m_schedule[hb->iterations & 1u] * mu_t;
  }
  hb->iterations += 1u;
}

__global__ void __launch_bounds__(256) optimizer_kernel(float* __restrict__ params, float* __restrict__ grads,
                                                        float* __restrict__ slot0, float* __restrict__ slot1,
                                                        int n4, int kind, HyperBlock* hb, int advance) {
  // advance = 0: an early bucket of the step (the dense layers, updated while the convolution backward still runs); the step
  // counter / Nadam schedule move on only with the step's last bucket
  const OptCoef c = opt_coefficients(hb, kind);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(params)[i];
    const float4 gv = reinterpret_cast<float4*>(grads)[i];
    float4 a = reinterpret_cast<float4*>(slot0)[i];
    float4 b = slot1 ? reinterpret_cast<float4*>(slot1)[i] : make_float4(0, 0, 0, 0);
    opt_update(kind, hb, c, gv.x, pv.x, a.x, b.x);
    opt_update(kind, hb, c, gv.y, pv.y, a.y, b.y);
    opt_update(kind, hb, c, gv.z, pv.z, a.z, b.z);
    opt_update(kind, hb, c, gv.w, pv.w, a.w, b.w);
    reinterpret_cast<float4*>(params)[i] = pv;
    reinterpret_cast<float4*>(slot0)[i] = a;
    if (slot1) reinterpret_cast<float4*>(slot1)[i] = b;
    reinterpret_cast<float4*>(grads)[i] = make_float4(0, 0, 0, 0);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&hb->ticket, 1u);
    if (tk == gridDim.x - 1) { hb->ticket = 0; if (advance) opt_advance_state(hb, kind); __threadfence(); }
  }
}

// =============================================================================================
// Batch assembly from the device-resident dataset: rows perm[cursor .. cursor+B) (clamped), sample weights
// 1 for real rows and 0 for the padding of a short last batch.
// =============================================================================================
struct GatherParams {
  float* xb; float* yb; float* sw;
  int row_f4;          // floats/4 per sample
  int label_w;         // floats per label row
  int B;               // batch rows
  const HyperBlock* hb;   // dataset pointers, row count and batch_cursor (first index into perm for this step)
};
__global__ void __launch_bounds__(256) gather_batch_kernel(const GatherParams p) {
  pdl_launch(); pdl_wait();
  const int b = blockIdx.y;
  const int cursor = (int)p.hb->batch_cursor;
  const int n_total = p.hb->n_total;
  const int* perm = p.hb->perm;
  int n_valid = n_total - cursor; if (n_valid > p.B) n_valid = p.B; if (n_valid < 1) n_valid = 1;
  const int src_b = b < n_valid ? b : n_valid - 1;
  int pos = cursor + src_b; if (pos >= n_total) pos = n_total - 1;
  const int row = perm ? perm[pos] : pos;
  const float4* src = reinterpret_cast<const float4*>(p.hb->data_x) + (size_t)row * p.row_f4;
  float4* dst = reinterpret_cast<float4*>(p.xb) + (size_t)b * p.row_f4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.row_f4; i += gridDim.x * blockDim.x) dst[i] = __ldg(src + i);
  if (blockIdx.x == 0) {
    for (int j = threadIdx.x; j < p.label_w; j += blockDim.x) p.yb[(size_t)b * p.label_w + j] = p.hb->data_y[(size_t)row * p.label_w + j];
    if (threadIdx.x == 0) p.sw[b] = b < n_valid ? 1.f : 0.f;
  }
}

__global__ void set_sample_weights_kernel(float* sw, int B, int n_valid) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) sw[b] = b < n_valid ? 1.f : 0.f;
}

}  // namespace cib
