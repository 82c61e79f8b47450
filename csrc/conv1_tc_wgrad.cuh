// Weight + bias gradient of the first convolution layer (rpv.py:46-49, mnist.py:49; op K9 of SURVEY 2.7.1) on the tensor cores (see conv1_tc.cuh for the producer design).
//
//   D[taps][CO] += A^T[32 taps][128 px] x dZ^T[CO][128 px]        K = the 128 pixels of a (16 w x 8 h) tile, 16 MMAs per tile
//
// Both operands are K-major (pixels contiguous along a 128-byte row, SWIZZLE_128B), stored as four column blocks of 32
// pixels: A^T block = [32 rows][128 B] (4 KB), dZ^T block = [CO rows][128 B].  The producers transpose while they write:
// every element is a 4-byte shared-memory store anyway.  (The MN-major formulation, which needs no transposition, ran at
// 160 cycles per MMA: tcgen05 reads MN-major tf32 operands at about 32 B/clk; K-major operands stream at full width.)
// Row 9*CI of A^T is all ones, so accumulator row 9*CI is the bias gradient.  dZ rows are rebuilt from the gradient at
// the pooled grid and the arg-max bits of the stored activation (or copied when the layer has no pooling).
// tcgen05 wants M = 128: rows 32..127 of an A block are whatever follows it in shared memory (finite: the stages are
// zero-initialised and only ever hold finite data) and the matching accumulator rows are never read.
// One accumulator for the CTA's whole tile slice; epilogue = red.global.add of 9*CI+1 rows.
#pragma once
#include "conv1_tc.cuh"

namespace cib {

__global__ void __launch_bounds__(C1_WG_THREADS, 1) conv1_tc_wgrad_kernel(const Conv1TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[8];
  __shared__ __align__(8) uint64_t empty_bar[8];
  __shared__ __align__(8) uint64_t done_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) float halo_s[C1_HALO_SLOTS][C1_HALO_FLOATS];

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int S = p.stages, CO = p.CO;
  const uint32_t zblk = (uint32_t)CO * 128u;                 // one 32-pixel column block of dZ^T
  const uint32_t stage_bytes = C1_TILE_BYTES + 4u * zblk;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
  // Four accumulators, used round-robin by the K steps: back-to-back tcgen05.mma into the SAME accumulator are serialised by
  // its read-modify-write latency (~100 cycles each for these tiny N = 16..64 MMAs, measured), independent ones pipeline.
  uint32_t tmem_cols = 32; while ((int)tmem_cols < 4 * CO) tmem_cols <<= 1;

  const int per = (p.num_tiles + gridDim.x - 1) / gridDim.x;
  const int t_begin = blockIdx.x * per;
  const int t_end = min(p.num_tiles, t_begin + per);
  const int ntile = t_end - t_begin;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full_bar[s]), C1_LOADERS / 32); mbar_init(smem_u32(&empty_bar[s]), 1); }
    mbar_init(smem_u32(&done_bar), 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  {
    // zero the stages and the slack behind them, then the all-ones row k = 9*CI of every A^T block
    float4* z4 = reinterpret_cast<float4*>(smem_al);
    const uint32_t n4 = (S * stage_bytes + C1_TILE_BYTES) / 16;
    for (uint32_t i = threadIdx.x; i < n4; i += blockDim.x) z4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    const int k1 = 9 * p.CI;
    for (int i = threadIdx.x; i < S * 128; i += blockDim.x) {
      const int s = i >> 7, m = i & 127;
      *reinterpret_cast<float*>(smem_al + (size_t)s * stage_bytes + (m >> 5) * 4096u + k1 * 128u +
                                (((((uint32_t)m & 31u) >> 2) ^ ((uint32_t)k1 & 7u)) << 4) + (m & 3) * 4u) = 1.0f;
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_smem, 0);

  if (ntile > 0) {
    if (warp == 0) {
      const bool leader = elect_one();
      const uint32_t idesc = make_idesc_tf32(128, CO, 0, 0);
      const uint64_t a_desc0 = make_sdesc(smem_base, 16, 1024, SWZ_128);
      const uint64_t z_desc0 = make_sdesc(smem_base + C1_TILE_BYTES, 16, 1024, SWZ_128);
      const uint32_t stage_u = stage_bytes >> 4, ablk_u = 4096u >> 4, zblk_u = zblk >> 4;
      for (int i = 0; i < ntile; ++i) {
        const int s = i % S; const uint32_t ph = (i / S) & 1;
        mbar_wait(smem_u32(&full_bar[s]), ph, 0x200 | s);
        tc_fence_after();
        const uint64_t a_st = a_desc0 + (uint64_t)(s * stage_u);
        const uint64_t z_st = z_desc0 + (uint64_t)(s * stage_u);
#pragma unroll
        for (int ks = 0; ks < 16; ++ks)
          if (leader) umma_tf32(tmem_base + (ks & 3) * CO, a_st + (uint64_t)((ks >> 2) * ablk_u + (ks & 3) * 2u),
                                z_st + (uint64_t)((ks >> 2) * zblk_u + (ks & 3) * 2u), idesc, (i > 0 || ks >= 4) ? 1u : 0u);
        if (leader) umma_commit(smem_u32(&empty_bar[s]));
        __syncwarp();
      }
      if (leader) umma_commit(smem_u32(&done_bar));
    } else if (warp >= 6) {
      const int lt = threadIdx.x - 192, lw = lt >> 5;
      C1Halo H; c1_make_halo(H, lt, p.CI);
      const int rowlen = 18 * p.CI;
      const uint32_t halo_base = smem_u32(&halo_s[0][0]);
      // A^T expansion: this warp owns the taps k = lw + 8 j; lane = pixel inside a 32-pixel column block
      uint32_t koff[4]; bool kok[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = lw + 8 * j;
        kok[j] = k < 9 * p.CI;
        const int tap = k / p.CI, ci = k - tap * p.CI, kh = tap / 3, kw = tap - kh * 3;
        koff[j] = (uint32_t)(kh * rowlen + kw * p.CI + ci) * 4u;
      }
      const uint32_t lane_src = (uint32_t)((lane >> 4) * rowlen + (lane & 15) * p.CI) * 4u;
      // dZ^T: items of 4 channels, prefetched one tile ahead.  pool: item = (pooled pixel of the 4 x 8 pooled tile, channel
      // group), <= 2 per thread; no pool: item = (pixel, channel group), CO / 8 per thread (<= 8)
      const int per_pix = CO / 4;
      const int Hs2 = p.Ho >> 1, Ws2 = p.Wo >> 1;
      const int z_items = p.pool ? 32 * per_pix : 128 * per_pix;
      uint4 gq[8], aq[2];
      auto fetch_z = [&](int i) {
        int n, h0, w0;
        if (i < ntile) c1_tile_coords(p, t_begin + i, n, h0, w0); else { n = 0; h0 = 1 << 20; w0 = 1 << 20; }
        if (p.pool) {
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            gq[u] = make_uint4(0, 0, 0, 0); aq[u] = gq[u];
            const int item = lt + u * C1_LOADERS;
            if (item < z_items) {
              const int pp = item / per_pix, c4 = item - pp * per_pix;
              const int gph = (h0 >> 1) + (pp >> 3), gpw = (w0 >> 1) + (pp & 7);
              if (gph < Hs2 && gpw < Ws2) {
                const size_t idx = (((size_t)n * Hs2 + gph) * Ws2 + gpw) * CO + c4 * 4;
                gq[u] = *reinterpret_cast<const uint4*>(p.g + idx);
                aq[u] = *reinterpret_cast<const uint4*>(p.act + idx);
              }
            }
          }
        } else {
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            gq[u] = make_uint4(0, 0, 0, 0);
            const int item = lt + u * C1_LOADERS;
            if (item < z_items) {
              const int m = item / per_pix, c4 = item - m * per_pix;
              const int gh = h0 + (m >> 4), gw = w0 + (m & 15);
              if (gh < p.Ho && gw < p.Wo)
                gq[u] = *reinterpret_cast<const uint4*>(p.g + (((size_t)n * p.Ho + gh) * p.Wo + gw) * CO + c4 * 4);
            }
          }
        }
      };
      auto z_addr = [&](uint32_t z_st, int m, int co) -> uint32_t {
        return z_st + (uint32_t)(m >> 5) * zblk + (uint32_t)co * 128u + (((((uint32_t)m & 31u) >> 2) ^ ((uint32_t)co & 7u)) << 4) + ((uint32_t)m & 3u) * 4u;
      };
      for (int i = 0; i < C1_HALO_AHEAD; ++i) {
        if (i < ntile) c1_issue_halo(H, p, t_begin + i, halo_base + (i % C1_HALO_SLOTS) * (C1_HALO_FLOATS * 4u));
        cp_async_commit();
      }
      fetch_z(0);
      for (int i = 0; i < ntile; ++i) {
        if (i + C1_HALO_AHEAD < ntile)
          c1_issue_halo(H, p, t_begin + i + C1_HALO_AHEAD, halo_base + ((i + C1_HALO_AHEAD) % C1_HALO_SLOTS) * (C1_HALO_FLOATS * 4u));
        cp_async_commit();
        cp_async_wait<C1_HALO_AHEAD>();
        loaders_bar();
        const int s = i % S; const uint32_t ph = (i / S) & 1;
        if (lane == 0) mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1, 0x100 | s);
        __syncwarp();
        const uint32_t a_st = smem_base + s * stage_bytes;
        const uint32_t z_st = a_st + C1_TILE_BYTES;
        if (!(p.debug & 8)) {
          const uint32_t src0 = halo_base + (i % C1_HALO_SLOTS) * (C1_HALO_FLOATS * 4u) + lane_src;
          const uint32_t dst_lane = (((uint32_t)lane >> 2) << 4) + ((uint32_t)lane & 3u) * 4u;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (kok[j]) {
              const uint32_t k = (uint32_t)(lw + 8 * j);
              const uint32_t drow = a_st + k * 128u + (dst_lane ^ ((k & 7u) << 4));
#pragma unroll
              for (int mb = 0; mb < 4; ++mb)
                sts_f32(drow + mb * 4096u, lds_f32(src0 + koff[j] + (uint32_t)(mb * 2 * rowlen) * 4u));
            }
          }
        }
        if (!(p.debug & 4)) {
          if (p.pool) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int item = lt + u * C1_LOADERS;
              if (item < z_items) {
                const int pp = item / per_pix, c4 = item - pp * per_pix;
                const int py = pp >> 3, px = pp & 7;
                const uint32_t gv[4] = {gq[u].x & ~3u, gq[u].y & ~3u, gq[u].z & ~3u, gq[u].w & ~3u};
                const uint32_t ps[4] = {aq[u].x & 3u, aq[u].y & 3u, aq[u].z & 3u, aq[u].w & 3u};
#pragma unroll
                for (int qq = 0; qq < 4; ++qq) {
                  const int m = (2 * py + (qq >> 1)) * 16 + 2 * px + (qq & 1);
#pragma unroll
                  for (int e = 0; e < 4; ++e)
                    sts_f32(z_addr(z_st, m, 4 * c4 + e), __uint_as_float(ps[e] == (uint32_t)qq ? gv[e] : 0u));
                }
              }
            }
          } else {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const int item = lt + u * C1_LOADERS;
              if (item < z_items) {
                const int m = item / per_pix, c4 = item - m * per_pix;
                const uint32_t gv[4] = {gq[u].x, gq[u].y, gq[u].z, gq[u].w};
#pragma unroll
                for (int e = 0; e < 4; ++e) sts_f32(z_addr(z_st, m, 4 * c4 + e), __uint_as_float(gv[e]));
              }
            }
          }
          fetch_z(i + 1);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&full_bar[s]));
      }
    } else if (warp == 4) {                                // warp 4 owns TMEM lanes 0..31
      // accumulator rows 0 .. 9*CI-1 = dW rows, row 9*CI = bias gradient (lane = row)
      mbar_wait(smem_u32(&done_bar), 0, 0x300);
      tc_fence_after();
      const int K = 9 * p.CI;
      for (int c0 = 0; c0 < CO; c0 += 16) {
        uint32_t v[16], t[16];
        tmem_ld16(tmem_base + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int a = 1; a < 4; ++a) {
          tmem_ld16(tmem_base + a * CO + c0, t);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(t[j]));
        }
        if (lane <= K) {
          float* dst = lane < K ? p.dW + (size_t)lane * CO + c0 : p.db + c0;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            red_add_v4_f32(dst + 4 * j, __uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
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
