// Dense-layer GEMMs on tcgen05 (kind::tf32, fp32 accumulate in TMEM), operands staged by TMA.
//
//   C[M,N] (+)= A[M,K] * B[K,N]        M tile = 128 rows, N tile = BN (multiple of 32, <= 256), K step = 32
//
// Operand layouts (chosen so that no tensor is ever transposed in global memory):
//   A K-major  : A stored [M][K]  (activations for forward / dgrad)        tensor map dims (K, M)
//   A MN-major : A stored [K][M]  (activations^T for wgrad: [batch][in])   tensor map dims (M, K)
//   B K-major  : B stored [N][K]  (dense kernel (in,out) seen by dgrad)    tensor map dims (K, N)
//   B MN-major : B stored [K][N]  (dense kernel (in,out) seen by forward)  tensor map dims (N, K)
// K-major tiles use the 128-byte swizzle (32 fp32 per row); MN-major tf32 tiles must use the 128-byte swizzle with
// 32-byte atoms (UMMA layout SWIZZLE_128B_BASE32B / TMA SWIZZLE_128B_ATOM_32B), measured: the plain 128B swizzle
// gives wrong products for MN-major tf32.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 = epilogue.
// Replaces the Keras Dense layer's MKL sgemm calls (reference rpv.py:55, mnist.py:54).
#pragma once
#include "ptx.cuh"

namespace cib {

enum { GEMM_EPI_STORE = 0, GEMM_EPI_REDADD = 1, GEMM_EPI_UNPOOL = 2, GEMM_EPI_MASK = 3 };

struct GemmTcParams {
  CUtensorMap tmA;
  CUtensorMap tmB;
  float* C;            // EPI_STORE / EPI_REDADD destination, row-major [M][ldc]
  int ldc;
  int M, N, K;         // N, K as padded in memory
  int BN;              // N tile
  int k_per_split;     // multiple of 32
  int splits;
  int stages;
  int epi;
  // EPI_UNPOOL: C is the gradient w.r.t. a flattened pooled activation [M][Hp*Wp*Cp]; it is masked by
  // the stored activation (non-zero <=> ReLU active and not dropped; 2 LSBs = arg-max position) and
  // scattered to the full-resolution conv gradient dz [M][Hf][Wf][Cp].  Requires Cp % BN == 0.
  const float* act_prev;
  float* dz_out;
  int Hp, Wp, Cp, Hf, Wf;
  float scale;
};

static constexpr int GEMM_A_STAGE_BYTES = 128 * 32 * 4;   // 16 KB

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(192, 1) gemm_tc_kernel(const __grid_constant__ GemmTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[8];
  __shared__ __align__(8) uint64_t empty_bar[8];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;

  pdl_launch();
  if (threadIdx.x == 0) ktrace(0);
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  const int S = p.stages;
  const int BN = p.BN;
  const uint32_t stage_bytes = GEMM_A_STAGE_BYTES + (uint32_t)BN * 128u;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;

  const int n_tiles = (p.N + BN - 1) / BN;
  int t = blockIdx.x;
  const int split = t % p.splits; t /= p.splits;
  const int n_t = t % n_tiles;
  const int m_t = t / n_tiles;
  const int m0 = m_t * 128, n0 = n_t * BN;
  const int Kpad = (p.K + 31) & ~31;
  const int k_begin = split * p.k_per_split;
  int k_end = k_begin + p.k_per_split; if (k_end > Kpad) k_end = Kpad;
  const int nk = (k_end - k_begin) / 32;
  if (nk <= 0) return;                               // uniform over the CTA

  uint32_t tmem_cols = 32; while ((int)tmem_cols < BN) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA); tma_prefetch_desc(&p.tmB);
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full_bar[s]), 1); mbar_init(smem_u32(&empty_bar[s]), 1); }
    mbar_init(smem_u32(&tmem_full_bar), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_smem, 0);
  if (threadIdx.x == 0) ktrace(1);
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < nk; ++i) {
        const int s = i % S; const uint32_t ph = (i / S) & 1;
        mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1, 0x100 | s);
        const uint32_t fb = smem_u32(&full_bar[s]);
        mbar_expect_tx(fb, stage_bytes);
        const uint32_t a_dst = smem_base + s * stage_bytes;
        const uint32_t b_dst = a_dst + GEMM_A_STAGE_BYTES;
        const int k0 = k_begin + i * 32;
        if (!A_MN) {
          tma_load_2d(a_dst, &p.tmA, fb, k0, m0);                       // box (32 k, 128 m)
        } else {
#pragma unroll
          for (int b = 0; b < 4; ++b) tma_load_2d(a_dst + b * 4096, &p.tmA, fb, m0 + 32 * b, k0);   // box (32 m, 32 k)
        }
        if (!B_MN) {
          tma_load_2d(b_dst, &p.tmB, fb, k0, n0);                       // box (32 k, BN n)
        } else {
          for (int b = 0; b < BN / 32; ++b) tma_load_2d(b_dst + b * 4096, &p.tmB, fb, n0 + 32 * b, k0);  // box (32 n, 32 k)
        }
        if (i == 0) ktrace(2);
      }
      ktrace(3);
    }
  } else if (warp == 1) {
    // warp-uniform loop (descriptors stay in uniform registers); one elected lane issues the MMAs
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_tf32(128, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
    const uint64_t a0 = A_MN ? make_sdesc(smem_base, 4096, 512, SWZ_128_B32) : make_sdesc(smem_base, 16, 1024, SWZ_128);
    const uint64_t b0 = B_MN ? make_sdesc(smem_base + GEMM_A_STAGE_BYTES, 4096, 512, SWZ_128_B32)
                             : make_sdesc(smem_base + GEMM_A_STAGE_BYTES, 16, 1024, SWZ_128);
    const uint32_t a_ks = A_MN ? (1024u >> 4) : 2u, b_ks = B_MN ? (1024u >> 4) : 2u;
    for (int i = 0; i < nk; ++i) {
      const int s = i % S; const uint32_t ph = (i / S) & 1;
      mbar_wait(smem_u32(&full_bar[s]), ph, 0x200 | s);
      if (leader && i == 0) ktrace(4);
      if (leader && i == nk - 1) ktrace(5);
      tc_fence_after();
      const uint64_t ad = a0 + (uint64_t)(s * (stage_bytes >> 4));
      const uint64_t bd = b0 + (uint64_t)(s * (stage_bytes >> 4));
#pragma unroll
      for (int ks = 0; ks < 4; ++ks)                                    // 4 x (K = 8 tf32)
        if (leader) umma_tf32(tmem_base, ad + a_ks * ks, bd + b_ks * ks, idesc, (i > 0 || ks > 0) ? 1u : 0u);
      if (leader) umma_commit(smem_u32(&empty_bar[s]));
      __syncwarp();
    }
    if (leader) { umma_commit(smem_u32(&tmem_full_bar)); ktrace(6); }
  } else {
    // ---------------- epilogue: TMEM -> registers -> global ----------------
    const int q = warp & 3;                            // TMEM lane quarter this warp may touch
    const int row = m0 + q * 32 + lane;
    mbar_wait(smem_u32(&tmem_full_bar), 0, 0x300);
    if (threadIdx.x == 64) ktrace(7);
    tc_fence_after();
    const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int c0 = 0; c0 < BN; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tlane + c0, v);
      tmem_ld_wait();
      const int col = n0 + c0;
      if (row < p.M && col < p.N) {
        if (p.epi == GEMM_EPI_STORE) {
          float4* dst = reinterpret_cast<float4*>(p.C + (size_t)row * p.ldc + col);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                 __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
        } else if (p.epi == GEMM_EPI_REDADD) {
          float* dst = p.C + (size_t)row * p.ldc + col;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            red_add_v4_f32(dst + 4 * j, __uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                           __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
        } else if (p.epi == GEMM_EPI_MASK) {
          // gradient w.r.t. a stored activation of the same shape: keep it where the activation is non-zero
          const float* a = p.act_prev + (size_t)row * p.ldc + col;
          float4* dst = reinterpret_cast<float4*>(p.dz_out + (size_t)row * p.ldc + col);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint4 ab = *reinterpret_cast<const uint4*>(a + 4 * j);
            dst[j] = make_float4((ab.x & ~3u) ? __uint_as_float(v[4 * j]) * p.scale : 0.f,
                                 (ab.y & ~3u) ? __uint_as_float(v[4 * j + 1]) * p.scale : 0.f,
                                 (ab.z & ~3u) ? __uint_as_float(v[4 * j + 2]) * p.scale : 0.f,
                                 (ab.w & ~3u) ? __uint_as_float(v[4 * j + 3]) * p.scale : 0.f);
          }
        } else {
          // unpool scatter: an N tile covers the channels [cb, cb + BN) of ONE pooled pixel (BN divides Cp; tiles narrower than Cp
          // put more CTAs on the machine: 128 x 4096 x 128 with Cp = 64 is only 64 CTAs at BN = Cp)
          const int pp = n0 / p.Cp, cb = n0 - pp * p.Cp;
          const int ph_ = pp / p.Wp, pw_ = pp % p.Wp;
          const float* a = p.act_prev + (((size_t)row * p.Hp + ph_) * p.Wp + pw_) * p.Cp + cb + c0;
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
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) {
            const int hh = 2 * ph_ + (qq >> 1), ww = 2 * pw_ + (qq & 1);
            float4* dst = reinterpret_cast<float4*>(p.dz_out + (((size_t)row * p.Hf + hh) * p.Wf + ww) * p.Cp + cb + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              dst[j] = make_float4(pos[4 * j] == qq ? g[4 * j] : 0.f, pos[4 * j + 1] == qq ? g[4 * j + 1] : 0.f,
                                   pos[4 * j + 2] == qq ? g[4 * j + 2] : 0.f, pos[4 * j + 3] == qq ? g[4 * j + 3] : 0.f);
          }
        }
      }
    }
  }
  if (threadIdx.x == 64) ktrace(8);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
  if (threadIdx.x == 0) ktrace(9);
}

}  // namespace cib
