// 3x3 convolution forward / dgrad, second-generation producer (the one the engine uses whenever the filter fits in
// shared memory).  Measured with ncu on B200: the TMA-per-tap producer of conv_tc.cuh is bound by TMA box-row
// throughput (~10 cycles per 64..128-byte row, 1.3 M rows for the 16->32 layer: 51 us, tensor pipe 2 % active).
// Here instead:
//   * the whole 3x3 filter is loaded ONCE per CTA by TMA and stays resident in shared memory;
//   * per (16 w x 8 h)-pixel tile, four loader warps read the (10 x 18)-pixel halo once from global memory with
//     16-byte loads and store it as THREE kw-shifted, hardware-swizzled copies [10 rows of 16 pixels][C];
//     the kh shift of a tap is then a whole number of 16-row groups, i.e. just a descriptor start-address offset
//     (multiples of 1024/2048 bytes keep the swizzle phase), so 9 taps need 3 copies instead of 9 tiles;
//   * K is pipelined in 32-channel chunks (one pipeline stage = one chunk of one tile);
//   * tcgen05.mma accumulates all taps and chunks of a tile in TMEM; two accumulator stages overlap the epilogue
//     (bias + ReLU + 2x2 max-pool with arg-max bits + dropout, or the dgrad mask / un-pool scatter).
// A lone warp retires a dependent instruction only every ~5-7 cycles (measured with ncu + in-kernel timestamps), so the
// three roles are written for few instructions per tile and run on many warps:
// Warp roles (576 threads): warp 0 = filter TMA + MMA issuer + TMEM owner (descriptors pre-computed, 9 taps unrolled),
// warp 1 spare, warps 2..9 = epilogue (two warps per TMEM lane quarter, alternating 16-column chunks),
// warps 10..17 = halo loaders (per-thread item tables in registers, cp.async, up to 3 stages in flight).
// Measured balance per tile of the 16->32 layer (in-kernel timestamps): loaders 1.0 us with eight warps (2.0 us with four: the
// cp.async rate is limited per WARP as well as per SM), MMA issue 0.9 us, epilogue 1.35 us.  A variant with sixteen epilogue
// warps in two tile-parity groups on four accumulator stages and four loader warps was correct but slower (the loaders became
// the bottleneck, and 672 threads cap the kernel at 80 registers).  A 28-warp variant with setmaxnreg rebalancing (producer
// warpgroups 40 registers, epilogue warpgroups 96; 11 loader + 16 epilogue warps) was correct as well but every role got slower
// (loaders 1.5 us, epilogue 1.6 us per tile; step 249 us against 228 us), so the roles are not independent: they share the SM's
// LSU / issue resources, and the way forward is less work per tile (shared->shared kw copies, pooling through shared memory),
// not more warps.
#pragma once
#include "ptx.cuh"
#include "conv_tc.cuh"

namespace cib {

struct ConvTc2Params {
  CUtensorMap tmB;       // filter [9*Cin_pad rows][Cout_pad]: fwd box (32, KK) ATOM_32B; dgrad box (cb, NN) SW128/SW64
  const float* src;      // NHWC source tensor [n][Hs][Ws][KK]
  int Hs, Ws;
  int b_mn;              // 1 = forward, 0 = dgrad
  int KK, NN;            // GEMM K (source channels: 16 or multiple of 32), N (result channels)
  int n_img, Ho, Wo;     // result pixel grid
  int tiles_w, tiles_h, num_tiles;
  int org_dh, org_dw;    // halo origin relative to the tile origin (min tap offsets)
  int tap_ih[9], tap_iw[9], tap_row[9];   // per tap: row-group shift (0..2), copy index (0..2), filter-matrix row
  int stages;            // A pipeline stages
  int epi;
  const float* bias; float* out; int pool; float drop_rate; unsigned int layer_id; const HyperBlock* hb;
  const float* act_prev; float* dz_out; int Hf, Wf; float scale;
};

static constexpr int C2_ROWS = 160;          // 10 halo rows x 16 pixels per kw copy
static constexpr int C2_THREADS = 576;
static constexpr int C2_LOADERS = 256;       // loader threads (warps 10..17)
static constexpr int C2_LD_WARP0 = 10;

__device__ __forceinline__ float4 ldg_nc_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void sts_f4(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// 16-byte asynchronous global->shared copy (LDGSTS); src_bytes = 0 zero-fills (used for padding / out-of-grid pixels)
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void c2_tile_coords(const ConvTc2Params& p, int tile, int& n, int& h0, int& w0) {
  const int tw = tile % p.tiles_w; tile /= p.tiles_w;
  const int th = tile % p.tiles_h; n = tile / p.tiles_h;
  w0 = tw * 16; h0 = th * 8;
}

__global__ void __maxnreg__(88) conv_tc2_kernel(const __grid_constant__ ConvTc2Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[4];
  __shared__ __align__(8) uint64_t empty_bar[4];
  __shared__ __align__(8) uint64_t tfull_bar[2];
  __shared__ __align__(8) uint64_t tempty_bar[2];
  __shared__ __align__(8) uint64_t b_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ float bias_s[256];

  pdl_launch();
  if (threadIdx.x == 0) ktrace(0);
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform for the compiler (uniform datapath)
  const int lane = threadIdx.x & 31;
  const int S = p.stages;
  const int KK = p.KK, NN = p.NN;
  const uint32_t RB = KK >= 32 ? 128u : 64u;               // bytes per pixel row inside one K chunk
  const int nchunks = KK >= 32 ? KK / 32 : 1;
  const int cpr = RB / 16;                                 // 16-byte pieces per row
  const uint32_t copy_bytes = C2_ROWS * RB;                // one kw copy of one K chunk
  const uint32_t stage_bytes = 3 * copy_bytes;
  const uint32_t b_bytes = 9u * KK * NN * 4u;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_base = smem_base;
  const uint32_t a_base = smem_base + ((b_bytes + 1023u) & ~1023u);

  uint32_t tmem_cols = 32; while ((int)tmem_cols < 2 * NN) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&p.tmB);
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&full_bar[s]), C2_LOADERS / 32); mbar_init(smem_u32(&empty_bar[s]), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(smem_u32(&tfull_bar[s]), 1); mbar_init(smem_u32(&tempty_bar[s]), 8); }
    mbar_init(smem_u32(&b_bar), 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_smem), tmem_cols);
  if (p.epi == CONV_EPI_FWD) for (int i = threadIdx.x; i < NN; i += blockDim.x) bias_s[i] = p.bias[i];   // parameters: constant within a step
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_smem, 0);
  if (threadIdx.x == 0) ktrace(1);
  // Programmatic dependent launch: everything above overlapped with the previous kernel's tail, and so does the FILTER load below
  // (warp 0 issues it before waiting: parameters are only written by the update kernels, which come after every convolution of the
  // step; the early bucket touches the dense stack only).  Each role calls pdl_wait() right before its first access to
  // activations / gradients.

  if (warp == 0) {
    // ====================== filter load (once) + MMA issue ======================
    // The whole warp runs this loop in warp-uniform control flow so that descriptors live in uniform registers
    // (issuing from inside an `if (lane == 0)` region cost 6 R2UR moves per tcgen05.mma); one elected lane issues.
    const bool leader = elect_one();
    {
      const uint32_t bb = smem_u32(&b_bar);
      if (leader) {
        mbar_expect_tx(bb, b_bytes);
        for (int tap = 0; tap < 9; ++tap) {
          const uint32_t dst = b_base + tap * (KK * NN * 4u);
          if (p.b_mn) { for (int b = 0; b < NN / 32; ++b) tma_load_2d(dst + b * (KK * 128u), &p.tmB, bb, b * 32, p.tap_row[tap]); }
          else        { for (int kc = 0; kc < nchunks; ++kc) tma_load_2d(dst + kc * (NN * RB), &p.tmB, bb, kc * (RB / 4), p.tap_row[tap]); }
        }
        ktrace(2);
      }
      __syncwarp();
      mbar_wait(bb, 0, 0x500);
      if (leader) ktrace(3);
      const uint32_t idesc = make_idesc_tf32(128, NN, 0, p.b_mn ? 1 : 0);
      const uint32_t a_layout = KK >= 32 ? SWZ_128 : SWZ_64;
      // Loop-invariant descriptor pieces (the address field counts 16-byte units, so offsets are plain adds).
      const uint64_t a_desc0 = make_sdesc(a_base, 16, 8 * RB, a_layout);
      const uint64_t b_desc0 = p.b_mn ? make_sdesc(b_base, KK * 128u, 512, SWZ_128_B32)
                             : (KK >= 32 ? make_sdesc(b_base, 16, 1024, SWZ_128) : make_sdesc(b_base, 16, 512, SWZ_64));
      const uint32_t stage_u = stage_bytes >> 4;
      const uint32_t tap_u = (KK * NN * 4u) >> 4;            // filter bytes per tap, 16-byte units
      const uint32_t b_kc = p.b_mn ? (4u * 1024u >> 4) : ((NN * 128u) >> 4);
      const uint32_t b_ks = p.b_mn ? (1024u >> 4) : 2u;
      int it = 0, ti = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++ti) {
        const int as = ti & 1; const uint32_t aph = (ti >> 1) & 1;
        mbar_wait(smem_u32(&tempty_bar[as]), aph ^ 1, 0x400 | as);
        tc_fence_after();
        const uint32_t tacc = tmem_base + as * NN;
        for (int kc = 0; kc < nchunks; ++kc, ++it) {
          const int s = it % S; const uint32_t ph = (it / S) & 1;
          mbar_wait(smem_u32(&full_bar[s]), ph, 0x200 | s);
          if (leader && it < 4) ktrace(4 + it);
          tc_fence_after();
          const uint64_t a_st = a_desc0 + (uint64_t)(s * stage_u);
          const uint64_t b_st = b_desc0 + (uint64_t)(kc * b_kc);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const uint64_t ad = a_st + (uint64_t)((p.tap_iw[tap] * copy_bytes + p.tap_ih[tap] * 16u * RB) >> 4);
            const uint64_t bd = b_st + (uint64_t)(tap * tap_u);
            if (KK >= 32) {
#pragma unroll
              for (int ks = 0; ks < 4; ++ks)
                if (leader) umma_tf32(tacc, ad + 2u * ks, bd + b_ks * ks, idesc, (kc > 0 || tap > 0 || ks > 0) ? 1u : 0u);
            } else {
#pragma unroll
              for (int ks = 0; ks < 2; ++ks)
                if (leader) umma_tf32(tacc, ad + 2u * ks, bd + b_ks * ks, idesc, (kc > 0 || tap > 0 || ks > 0) ? 1u : 0u);
            }
          }
          if (leader) umma_commit(smem_u32(&empty_bar[s]));
        }
        if (leader) umma_commit(smem_u32(&tfull_bar[as]));
        if (leader && ti < 2) ktrace(8 + ti);
        __syncwarp();
      }
      if (leader) ktrace(10);
    }
  } else if (warp >= C2_LD_WARP0) {
    // ====================== halo loaders: 8 warps, cp.async with up to 3 stages in flight ======================
    const int lt = threadIdx.x - C2_LD_WARP0 * 32;          // 0..255
    const int items = 180 * cpr;                            // (halo pixel, 16-byte piece) pairs per K chunk
    const int lag = S >= 3 ? 2 : (S - 1);                   // stages whose copies may still be in flight
    // Tile-invariant description of this thread's items (<= 6): source offset from the halo origin (floats), halo
    // coordinates, and the swizzled destination of each kw copy (0xFFFFFFFF = this pixel is not part of that copy).
    constexpr int MAXI = 6;
    int src_off[MAXI], hhw[MAXI]; uint32_t dst[MAXI][3];
#pragma unroll
    for (int k = 0; k < MAXI; ++k) {
      const int i = lt + k * C2_LOADERS;
      src_off[k] = 0; hhw[k] = -1; dst[k][0] = dst[k][1] = dst[k][2] = 0xFFFFFFFFu;
      if (i < items) {
        const int pix = i / cpr, c = i - pix * cpr;
        const int hh = pix / 18, hw = pix - hh * 18;
        src_off[k] = (hh * p.Ws + hw) * KK + c * 4;
        hhw[k] = hh | (hw << 8);
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const int ww = hw - kw;
          if (ww >= 0 && ww < 16) {
            const int row = hh * 16 + ww;
            const int sw = KK >= 32 ? (c ^ (row & 7)) : (c ^ ((row >> 1) & 3));
            dst[k][kw] = kw * copy_bytes + row * RB + sw * 16;
          }
        }
      }
    }
    int it = 0;
    pdl_wait();                                             // the item tables above are tile-invariant arithmetic
    auto publish = [&](int j) {                             // copies of pipeline step j have landed: hand the stage to the MMA warp
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&full_bar[j % S]));
      if (lt == 0 && j < 3) ktrace(13 + 2 * j);
    };
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      int n, h0, w0; c2_tile_coords(p, tile, n, h0, w0);
      const int gh0 = h0 + p.org_dh, gw0 = w0 + p.org_dw;
      const float* img = p.src + (size_t)n * p.Hs * p.Ws * KK;
      const float* org = img + ((long long)gh0 * p.Ws + gw0) * KK;   // may point before the image at the border: only
                                                                      // dereferenced for in-bounds pixels
      for (int kc = 0; kc < nchunks; ++kc, ++it) {
        const int s = it % S; const uint32_t ph = (it / S) & 1;
        if (lane == 0) mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1, 0x100 | s);
        __syncwarp();
        if (lt == 0 && it < 3) ktrace(12 + 2 * it);
        const uint32_t st_base = a_base + s * stage_bytes;
#pragma unroll
        for (int k = 0; k < MAXI; ++k) {
          if (hhw[k] >= 0) {
            const int gh = gh0 + (hhw[k] & 255), gw = gw0 + (hhw[k] >> 8);
            const bool ok = (unsigned)gh < (unsigned)p.Hs && (unsigned)gw < (unsigned)p.Ws;
            const float* src = ok ? org + src_off[k] + kc * 32 : img;
            const uint32_t nb = ok ? 16u : 0u;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw)
              if (dst[k][kw] != 0xFFFFFFFFu) cp_async16(st_base + dst[k][kw], src, nb);
          }
        }
        cp_async_commit();
        if (it >= lag) {
          if (lag == 2) cp_async_wait<2>(); else if (lag == 1) cp_async_wait<1>(); else cp_async_wait<0>();
          publish(it - lag);
        }
      }
    }
    // drain
    if (lag == 2 && it >= 2) { cp_async_wait<1>(); publish(it - 2); }
    if (lag >= 1 && it >= 1) { cp_async_wait<0>(); publish(it - 1); }
    if (lt == 0) ktrace(18);
  } else if (warp >= 2) {
    // ====================== epilogue warps (2..9): two per TMEM lane quarter ======================
    pdl_wait();
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;                       // 0: chunks 0,2,4..   1: chunks 1,3,5..                                 // TMEM lane quarter; rows q*32 .. q*32+31 = tile rows hh = 2q, 2q+1
    const int lh = lane >> 4, lw = lane & 15;
    const unsigned int step = p.hb ? p.hb->iterations : 0u;
    const unsigned int seed = p.hb ? p.hb->seed : 0u;
    const bool do_drop = p.drop_rate > 0.f && p.hb && p.hb->training;
    const float keep_scale = do_drop ? 1.0f / (1.0f - p.drop_rate) : 1.0f;
    const int gi = (lane & 1) + 2 * lh;                     // position inside the 2x2 pooling window
    int ti = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++ti) {
      const int as = ti & 1; const uint32_t aph = (ti >> 1) & 1;
      int n, h0, w0; c2_tile_coords(p, tile, n, h0, w0);
      const int h = h0 + 2 * q + lh, w = w0 + lw;
      const bool inb = h < p.Ho && w < p.Wo;
      // dgrad: fetch the first chunk of the stored activation (mask + arg-max bits) before waiting for the accumulator
      const float* act_row = (p.epi != CONV_EPI_FWD && inb) ? p.act_prev + (((size_t)n * p.Ho + h) * p.Wo + w) * NN : nullptr;
      uint4 ab_next[4];
      if (act_row) {
#pragma unroll
        for (int j = 0; j < 4; ++j) ab_next[j] = *reinterpret_cast<const uint4*>(act_row + 16 * half + 4 * j);
      }
      mbar_wait(smem_u32(&tfull_bar[as]), aph, 0x300 | as);
      if (threadIdx.x == 64 && ti < 3) ktrace(20 + 2 * ti);
      tc_fence_after();
      const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16) + as * NN;
      for (int c0 = 16 * half; c0 < NN; c0 += 32) {
        uint32_t v[16];
        tmem_ld16(tlane + c0, v);
        tmem_ld_wait();
        if (p.epi == CONV_EPI_FWD) {
          if (p.pool) {
            const int Hs2 = p.Ho >> 1, Ws2 = p.Wo >> 1;
            // Stage the warp-synchronous ops so that the 16 channels' shuffles issue back to back (a per-channel
            // dependent chain of SHFL/VOTE serialises: measured 3.6 k cycles per chunk before this restructuring).
            const int wbase = lane & ~17;                     // first lane of my 2x2 window (lanes +0, +1, +16, +17)
            float x[16], m[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = fmaxf(__uint_as_float(v[j]) + bias_s[c0 + j], 0.f);
#pragma unroll
            for (int j = 0; j < 16; ++j) m[j] = fmaxf(x[j], __shfl_xor_sync(0xffffffffu, x[j], 1));
#pragma unroll
            for (int j = 0; j < 16; ++j) m[j] = fmaxf(m[j], __shfl_xor_sync(0xffffffffu, m[j], 16));
            uint32_t eq[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) eq[j] = __ballot_sync(0xffffffffu, x[j] == m[j]) >> wbase;   // who holds the maximum
            // lane `gi` of the window stores channels 4*gi .. 4*gi+3: pick them with selects (no divergent branches)
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float mv = m[e]; uint32_t qv = eq[e];
              mv = gi == 1 ? m[4 + e] : mv;  qv = gi == 1 ? eq[4 + e] : qv;
              mv = gi == 2 ? m[8 + e] : mv;  qv = gi == 2 ? eq[8 + e] : qv;
              mv = gi == 3 ? m[12 + e] : mv; qv = gi == 3 ? eq[12 + e] : qv;
              const uint32_t w4 = (qv & 3u) | ((qv >> 14) & 12u);                  // bits: pos 0,1 (row 0), 2,3 (row 1)
              const uint32_t pv = (uint32_t)__ffs((int)w4) - 1u;                   // first (lowest) position wins ties
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
        } else if (inb) {
          uint4 ab[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) ab[j] = ab_next[j];
          if (c0 + 32 < NN) {
#pragma unroll
            for (int j = 0; j < 4; ++j) ab_next[j] = *reinterpret_cast<const uint4*>(act_row + c0 + 32 + 4 * j);
          }
          if (p.epi == CONV_EPI_DGRAD_MASK) {
            float4* dst = reinterpret_cast<float4*>(p.dz_out + (((size_t)n * p.Ho + h) * p.Wo + w) * NN + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              dst[j] = make_float4((ab[j].x & ~3u) ? __uint_as_float(v[4 * j]) * p.scale : 0.f,
                                   (ab[j].y & ~3u) ? __uint_as_float(v[4 * j + 1]) * p.scale : 0.f,
                                   (ab[j].z & ~3u) ? __uint_as_float(v[4 * j + 2]) * p.scale : 0.f,
                                   (ab[j].w & ~3u) ? __uint_as_float(v[4 * j + 3]) * p.scale : 0.f);
          } else {
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
              const int hh = 2 * h + (qq >> 1), ww = 2 * w + (qq & 1);
              float4* dst = reinterpret_cast<float4*>(p.dz_out + (((size_t)n * p.Hf + hh) * p.Wf + ww) * NN + c0);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint32_t bx[4] = {ab[j].x, ab[j].y, ab[j].z, ab[j].w};
                float y[4];
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  y[e] = ((bx[e] & ~3u) && (int)(bx[e] & 3u) == qq) ? __uint_as_float(v[4 * j + e]) * p.scale : 0.f;
                dst[j] = make_float4(y[0], y[1], y[2], y[3]);
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[as]));
      if (threadIdx.x == 64 && ti < 3) ktrace(21 + 2 * ti);
    }
    if (threadIdx.x == 64) ktrace(27);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, tmem_cols);
  if (threadIdx.x == 0) ktrace(28);
}

}  // namespace cib
