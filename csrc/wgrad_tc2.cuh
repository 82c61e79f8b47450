// 3x3 convolution weight + bias gradient (backward of Conv2D, rpv.py:46-48 / mnist.py:50; op K9 of SURVEY 2.7.1), second-generation producer (see conv_tc2.cuh for the measurement that
// motivated it: TMA box rows are too expensive for pixel-sized rows).
//
//   dW[kh][kw][ci][co] = sum_pix X[pix + (kh,kw) - pad][ci] * dZ[pix][co]         db[co] = sum_pix dZ[pix][co]
//
// GEMM view per (16 w x 8 h)-pixel tile, K = 128 pixels in 16 steps of 8:
//   A (M side, MN-major)  = the three kw-shifted halo copies of X stacked along M: M block (kw, ci-block of 32);
//                           the kh shift is a 16-row start offset of the descriptor, so ONE tcgen05.mma per kh
//                           produces all three kw taps:  D_kh[(kw,ci)][co]
//   B (N side, MN-major)  = the dZ tile [128 pixels][CO]
//   bias                  = one more MMA per K step whose A operand is a tile of ones: D_b[*][co] = sum_pix dZ[pix][co]
// Accumulators for the CTA's whole pixel slice stay in TMEM ((3*groups + 1) * CO columns <= 512); the epilogue adds the
// partial sums into the gradient buffer with red.global.add (co contiguous).
// MN-major tf32 operands use the 128-byte swizzle with 32-byte atoms; 16-channel inputs are stored as 128-byte rows
// whose upper half stays zero.
// Warp roles (448 threads): warp 0 = MMA issuer + TMEM owner (descriptors pre-computed), warps 2..5 = epilogue,
// warps 6..13 = loaders (cp.async, up to 3 tiles in flight).
#pragma once
#include "ptx.cuh"
#include "conv_tc2.cuh"

namespace cib {

struct WgradTc2Params {
  const float* x;       // [n][Hi][Wi][ci_mem]
  const float* dz;      // [n][Ho][Wo][CO]
  float* dW;            // [9][ci_mem][CO]
  float* db;            // [CO]
  int Hi, Wi, ci_mem;   // ci_mem = 16 or a multiple of 32
  int n_img, Ho, Wo, CO;
  int pad;
  int tiles_w, tiles_h, num_tiles;
  int stages;
};

__device__ __forceinline__ int swz_b32(int piece16, int row) {      // 16-byte piece index (0..7) inside a 128-byte row
  return ((((piece16 >> 1) ^ (row & 3)) << 1) | (piece16 & 1));
}

static constexpr int W2_THREADS = 448;       // warp 0 MMA, 1 spare, 2..5 epilogue, 6..13 loaders
static constexpr int W2_LOADERS = 256;

__global__ void __launch_bounds__(W2_THREADS, 1) wgrad_tc2_kernel(const WgradTc2Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[4];
  __shared__ __align__(8) uint64_t empty_bar[4];
  __shared__ __align__(8) uint64_t done_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ int4 pix_lut[180];      // per halo pixel: {source pixel offset, hh | hw << 8, row of the kw = 0 copy, 0}

  pdl_launch();
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  const int S = p.stages;
  const int CO = p.CO;
  const int cib = p.ci_mem >= 32 ? p.ci_mem / 32 : 1;      // 32-channel blocks of X
  const int cob = CO / 32;
  const int nmb = 3 * cib;                                  // M blocks (kw, ci block)
  const int ngroups = (nmb + 3) / 4;
  const uint32_t reg_bytes = C2_ROWS * 128u;                // one (kw, ci-block) copy: 160 rows x 128 B
  const uint32_t a_bytes = nmb * reg_bytes;
  const uint32_t z_bytes = cob * 128u * 128u;               // dZ tile: [co block][128 rows][128 B]
  const uint32_t stage_bytes = a_bytes + z_bytes;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ones_base = smem_base + S * stage_bytes;   // 4 blocks x 8 rows x 128 B of 1.0f, then slack
  const int acc_cols = (3 * ngroups + 1) * CO;
  uint32_t tmem_cols = 32; while ((int)tmem_cols < acc_cols) tmem_cols <<= 1;

  // contiguous slice of tiles for this CTA
  const int per = (p.num_tiles + gridDim.x - 1) / gridDim.x;
  const int t_begin = blockIdx.x * per;
  const int t_end = min(p.num_tiles, t_begin + per);
  const int ntile = t_end - t_begin;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full_bar[s]), W2_LOADERS / 32); mbar_init(smem_u32(&empty_bar[s]), 1); }
    mbar_init(smem_u32(&done_bar), 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  {
    // zero the A/B stages once (the upper half of 16-channel rows is never written again) and fill the ones tile
    float4* z4 = reinterpret_cast<float4*>(smem_raw + (smem_base - smem_u32(smem_raw)));
    const int n4 = (S * stage_bytes) / 16;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    float* ones = reinterpret_cast<float*>(smem_raw + (ones_base - smem_u32(smem_raw)));
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) ones[i] = 1.0f;
    for (int i = threadIdx.x; i < 180; i += blockDim.x) {
      const int hh = i / 18, hw = i - hh * 18;
      pix_lut[i] = make_int4(hh * p.Wi + hw, hh | (hw << 8), hh * 16 + hw, 0);
    }
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_smem, 0);
  pdl_wait();                                // the shared-memory set-up above (190 KB of zero fill) overlapped the previous kernel's tail

  if (ntile > 0) {
    if (warp == 0) {
      const bool leader = elect_one();       // warp-uniform loop, one elected lane issues (keeps descriptors uniform)
      {
        const uint32_t idesc = make_idesc_tf32(128, CO, 1, 1);
        // descriptors relative to stage 0; per stage / K step only the 16-byte-unit address field moves
        const uint64_t z_desc0 = make_sdesc(smem_base + a_bytes, 128u * 128u, 512, SWZ_128_B32);
        const uint64_t a_desc0 = make_sdesc(smem_base, reg_bytes, 512, SWZ_128_B32);
        const uint64_t o_desc = make_sdesc(ones_base, 1024, 512, SWZ_128_B32);
        const uint32_t stage_u = stage_bytes >> 4, kh_u = (16u * 128u) >> 4, g_u = (4u * reg_bytes) >> 4, ks_u = 1024u >> 4;
        const uint32_t bias_col = tmem_base + 3 * ngroups * CO;
        for (int i = 0; i < ntile; ++i) {
          const int s = i % S; const uint32_t ph = (i / S) & 1;
          mbar_wait(smem_u32(&full_bar[s]), ph, 0x200 | s);
          tc_fence_after();
          const uint64_t a_st = a_desc0 + (uint64_t)(s * stage_u);
          const uint64_t z_st = z_desc0 + (uint64_t)(s * stage_u);
#pragma unroll 4
          for (int ks = 0; ks < 16; ++ks) {
            const uint32_t acc = (i > 0 || ks > 0) ? 1u : 0u;
            const uint64_t bd = z_st + ks * ks_u;
            const uint64_t ak = a_st + ks * ks_u;
            if (ngroups == 1) {
#pragma unroll
              for (int kh = 0; kh < 3; ++kh) if (leader) umma_tf32(tmem_base + kh * CO, ak + kh * kh_u, bd, idesc, acc);
            } else {
              for (int kh = 0; kh < 3; ++kh)
                for (int g = 0; g < ngroups; ++g)
                  if (leader) umma_tf32(tmem_base + (kh * ngroups + g) * CO, ak + kh * kh_u + g * g_u, bd, idesc, acc);
            }
            if (leader) umma_tf32(bias_col, o_desc, bd, idesc, acc);
          }
          if (leader) umma_commit(smem_u32(&empty_bar[s]));
          __syncwarp();
        }
        if (leader) umma_commit(smem_u32(&done_bar));
      }
    } else if (warp >= 6) {
      const int lt = threadIdx.x - 192;                    // 0..255
      const int cpr_x = p.ci_mem >= 32 ? 8 : 4;            // 16-byte pieces of X per pixel and ci block
      const int per_pix_x = cpr_x * cib;
      const int x_items = 180 * per_pix_x;
      const int per_pix_z = CO / 4;
      const int z_items = 128 * per_pix_z;
      const int lag = S >= 3 ? 2 : (S - 1);
      auto publish = [&](int j) {
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&full_bar[j % S]));
      };
      for (int i = 0; i < ntile; ++i) {
        const int s = i % S; const uint32_t ph = (i / S) & 1;
        int tile = t_begin + i;
        const int tw = tile % p.tiles_w; tile /= p.tiles_w;
        const int th = tile % p.tiles_h; const int n = tile / p.tiles_h;
        const int w0 = tw * 16, h0 = th * 8;
        if (lane == 0) mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1, 0x100 | s);
        __syncwarp();
        const uint32_t a_dst = smem_base + s * stage_bytes;
        const uint32_t z_dst = a_dst + a_bytes;
        // ---- X halo -> three kw copies per ci block (cp.async, zero-fill outside the image)
        const float* ximg = p.x + (size_t)n * p.Hi * p.Wi * p.ci_mem;
        const long long pix_base = (long long)(h0 - p.pad) * p.Wi + (w0 - p.pad);
        const int xsh = 31 - __clz(per_pix_x);               // per_pix_x is 4, 8, 16, ... (power of two) -- see launcher
#pragma unroll 4
        for (int it = lt; it < x_items; it += W2_LOADERS) {
          const int pix = it >> xsh, c = it & (per_pix_x - 1);
          const int4 L = pix_lut[pix];
          const int hw = L.y >> 8;
          const int gh = h0 - p.pad + (L.y & 255), gw = w0 - p.pad + hw;
          const bool ok = (unsigned)gh < (unsigned)p.Hi && (unsigned)gw < (unsigned)p.Wi;
          const float* src = ok ? ximg + (pix_base + L.x) * p.ci_mem + c * 4 : ximg;
          const uint32_t nb = ok ? 16u : 0u;
          const int cb = c >> 3, piece = c & 7;              // (16-channel inputs: cb = 0, piece 0..3 = lower half row)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) {
            if ((unsigned)(hw - kw) < 16u) {
              const int row = L.z - kw;
              cp_async16(a_dst + (kw * cib + cb) * reg_bytes + row * 128u + swz_b32(piece, row) * 16u, src, nb);
            }
          }
        }
        // ---- dZ tile
        const float* zimg = p.dz + (size_t)n * p.Ho * p.Wo * CO;
        const int zsh = 31 - __clz(per_pix_z);
        const bool z_pow2 = (per_pix_z & (per_pix_z - 1)) == 0;
#pragma unroll 4
        for (int it = lt; it < z_items; it += W2_LOADERS) {
          const int pix = z_pow2 ? (it >> zsh) : (it / per_pix_z);
          const int c = it - pix * per_pix_z;
          const int gh = h0 + (pix >> 4), gw = w0 + (pix & 15);
          const bool ok = gh < p.Ho && gw < p.Wo;
          const float* src = ok ? zimg + ((size_t)gh * p.Wo + gw) * CO + c * 4 : zimg;
          const int cb = c >> 3, piece = c & 7;
          cp_async16(z_dst + cb * (128u * 128u) + pix * 128u + swz_b32(piece, pix) * 16u, src, ok ? 16u : 0u);
        }
        cp_async_commit();
        if (i >= lag) {
          if (lag == 2) cp_async_wait<2>(); else if (lag == 1) cp_async_wait<1>(); else cp_async_wait<0>();
          publish(i - lag);
        }
      }
      if (lag == 2 && ntile >= 2) { cp_async_wait<1>(); publish(ntile - 2); }
      if (lag >= 1 && ntile >= 1) { cp_async_wait<0>(); publish(ntile - 1); }
    } else if (warp >= 2) {
      const int q = warp & 3;
      mbar_wait(smem_u32(&done_bar), 0, 0x300);
      tc_fence_after();
      const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
      const int ci_out = p.ci_mem;
      for (int kh = 0; kh < 3; ++kh)
        for (int g = 0; g < ngroups; ++g) {
          const int mb = 4 * g + q;
          const int kw = mb / cib, cb = mb - kw * cib;
          const int ci = cb * 32 + lane;
          const bool ok = mb < nmb && ci < ci_out;
          for (int c0 = 0; c0 < CO; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tlane + (kh * ngroups + g) * CO + c0, v);
            tmem_ld_wait();
            if (ok) {
              float* dst = p.dW + ((size_t)((kh * 3 + kw) * ci_out + ci)) * CO + c0;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                red_add_v4_f32(dst + 4 * j, __uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                               __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
            }
          }
        }
      for (int c0 = 0; c0 < CO; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tlane + 3 * ngroups * CO + c0, v);
        tmem_ld_wait();
        if (q == 0 && lane == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            red_add_v4_f32(p.db + c0 + 4 * j, __uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                           __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, tmem_cols);
}

}  // namespace cib
