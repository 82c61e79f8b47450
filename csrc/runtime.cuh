// Host-side runtime helpers: error reporting, tensor-map construction, kernel launchers.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <mutex>
#include <nvtx3/nvToolsExt.h>     // header-only; the ranges are no-ops unless a tool (nsys, ncu --nvtx) is attached

#include "ptx.cuh"
#include "gemm_tc.cuh"
#include "conv_tc.cuh"
#include "wgrad_tc.cuh"
#include "conv_tc2.cuh"
#include "wgrad_tc2.cuh"
#include "conv1_tc.cuh"
#include "kernels_basic.cuh"
#include "allreduce.cuh"
#include "../include/cib200.h"

namespace cib {

static thread_local std::string g_last_error;
// Several trials can live in one process (an engine with --slots k runs them in threads).  Stream capture of one thread is
// invalidated by a device-wide synchronisation issued from any other thread (cudaDeviceSynchronize, cudaFree), so captures,
// (de)allocations and the Python side's dataset uploads all run inside this process-wide exclusive section.
static std::recursive_mutex g_exclusive_mu;
static unsigned int* g_err_mailbox_host = nullptr;      // reserved for future host-mapped mailbox
static int g_num_sms = 148;
static int g_device = -1;
// 1 (default): the dense stack's gradients are all-reduced and its parameters updated on a third stream while the convolution
// backward still runs (two buckets per step: reference Horovod reduces each gradient while backward continues, rpv.py:63-65);
// the early kernel is the 128-thread push all-reduce, which co-resides with the convolution CTAs instead of displacing them.
// 1 = under data parallelism only, 2 = always, 0 = one update at the end of the step.  CIB_EARLY_UPDATE / cib_set_early_update.
static int g_early_update = 1;
// 1: the convolution wgrad kernels run on the main stream between the dgrad kernels, chained by programmatic dependent launch, instead
// of on the side stream: they cannot share an SM with the dgrad kernels anyway (~200 KB of shared memory each).  CIB_SERIAL_BWD
static int g_serial_bwd = 0;
static int g_opt_grid_mult = 4;           // optimizer grid = at most this many blocks per SM (every block ends with one ticket atomic); CIB_OPT_GRID_MULT
static int g_copy_chunks = 4;             // pieces (= streams) a host-fed step's input copy is split into; CIB_COPY_CHUNKS (1..8)
static int g_head_spc = 8;                // samples per CTA of the multi-CTA loss head; CIB_HEAD_SPC
static int g_gemm_auto_bn = 1;            // CIB_GEMM_AUTOBN
static int g_dgrad_bn = 0;                // CIB_DGRAD_BN: force the N tile of the dense dgrad's un-pool epilogue (experiments)
static int g_fuse_dense_act = 1;          // 1: the last hidden layer's bias + ReLU + dropout runs inside the loss head (no separate kernel); CIB_FUSE_DENSE_ACT
static int g_head_version = 2;            // 2 = multi-CTA loss head (head_mc_kernel), 1 = single-CTA deterministic head
static int g_conv_version = 2;            // 2 = halo-copy producer (conv_tc2.cuh) when the filter fits, 1 = TMA-per-tap

// RAII NVTX range for the host-side phases (graph capture, step replay, host-fed steps, collectives set-up)
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

static int fail(const std::string& msg) { g_last_error = msg; return 1; }

#define CIB_CUDA(expr)                                                                                         \
  do {                                                                                                         \
    cudaError_t _e = (expr);                                                                                   \
    if (_e != cudaSuccess) {                                                                                   \
      return cib::fail(std::string(#expr) + " -> " + cudaGetErrorString(_e) + " @" + __FILE__ + ":" +         \
                       std::to_string(__LINE__));                                                              \
    }                                                                                                          \
  } while (0)
#define CIB_TRY(expr)            \
  do {                           \
    int _r = (expr);             \
    if (_r) return _r;           \
  } while (0)

// ---- kernel launch with optional programmatic dependent launch (see pdl_launch / pdl_wait in ptx.cuh) -----------------------
// The engine sets t_pdl right before it enqueues a kernel on its main stream whose stream predecessor is also a kernel; the launcher
// that finally issues the launch consumes the flag.  Only kernels that contain pdl_wait() are ever launched through launch_k.
static int g_use_pdl = 1;                       // CIB_PDL=0 / cib_set_pdl(0): ordinary stream serialization everywhere (A/B runs)
static thread_local bool t_pdl = false;
template <typename... KArgs, typename... Args>
static cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof cfg);
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  const bool pdl = t_pdl && g_use_pdl; t_pdl = false;
  if (pdl) {
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
  }
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace cib
#include "nvls_drv.cuh"
namespace cib {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled g_encode = nullptr;

static int ensure_driver() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CIB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess) return fail("cuTensorMapEncodeTiled not available from the driver");
  g_encode = reinterpret_cast<PFN_encodeTiled>(fn);
  return 0;
}

// swizzle_bytes: 128 or 64 (inner box extent in bytes must equal it), or SWZ_TMA_MN for MN-major tf32 operands
// (128-byte rows, 32-byte swizzle atoms)
static constexpr int SWZ_TMA_MN = 1232;
static int make_tmap(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box, int swizzle_bytes, const uint32_t* elem_strides = nullptr) {
  CIB_TRY(ensure_driver());
  cuuint64_t gd[5]; cuuint64_t gs[4]; cuuint32_t bx[5]; cuuint32_t es[5];
  // traversal stride s in a dimension: the box SPANS box[i] source elements and every s-th one is loaded (box[i] / s land in smem)
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = elem_strides ? elem_strides[i] : 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUtensorMapSwizzle sw = swizzle_bytes == SWZ_TMA_MN ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                        : swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                        : swizzle_bytes == 64  ? CU_TENSOR_MAP_SWIZZLE_64B
                        : swizzle_bytes == 32  ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled failed (%d): rank %d dims %llu %llu box %u %u swz %d", (int)r, rank,
             (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0), box[0], rank > 1 ? box[1] : 0, swizzle_bytes);
    return fail(buf);
  }
  return 0;
}
static int make_tmap_2d(CUtensorMap* tm, const void* base, uint64_t d0, uint64_t d1, uint64_t pitch_floats, uint32_t b0, uint32_t b1, int swz) {
  const uint64_t dims[2] = {d0, d1}; const uint64_t st[1] = {pitch_floats * 4}; const uint32_t box[2] = {b0, b1};
  return make_tmap(tm, base, 2, dims, st, box, swz);
}
// NHWC tensor [n][H][W][C] seen as dims (C, W, H, n), box (cb, 16, 2, 1)
static int make_tmap_nhwc(CUtensorMap* tm, const void* base, int n, int H, int W, int C, int cb, int swz = 0, int pix_stride = 1) {
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)n};
  const uint64_t st[3] = {(uint64_t)C * 4, (uint64_t)W * C * 4, (uint64_t)H * W * C * 4};
  const uint32_t box[4] = {(uint32_t)cb, 16u * pix_stride, 2u * pix_stride, 1};        // always 16 x 2 pixels in shared memory
  const uint32_t es[4] = {1, (uint32_t)pix_stride, (uint32_t)pix_stride, 1};
  return make_tmap(tm, base, 4, dims, st, box, swz ? swz : cb * 4, es);
}

// Dynamic shared memory opt-in is a per-FUNCTION attribute, so it is raised once to the hardware maximum for every
// tcgen05 kernel (init_kernel_attributes) and never lowered: a CUDA graph may hold nodes of the same function with
// different sizes, and tools that re-launch single graph nodes (ncu) read the attribute at that moment.
static constexpr int CIB_MAX_DYN_SMEM = 226 * 1024;      // 227 KB opt-in limit minus room for static shared memory
template <typename K>
static int set_smem(K kernel, size_t bytes) {
  if (bytes > (size_t)CIB_MAX_DYN_SMEM) return fail("kernel needs more dynamic shared memory than the SM has");
  (void)kernel;
  return 0;
}
template <typename K>
static int raise_smem(K kernel) {
  cudaFuncAttributes fa;
  CIB_CUDA(cudaFuncGetAttributes(&fa, kernel));
  int optin = 0, dev = 0;
  CIB_CUDA(cudaGetDevice(&dev));
  CIB_CUDA(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const int dyn = std::min<int>(CIB_MAX_DYN_SMEM, optin - (int)fa.sharedSizeBytes);
  CIB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// GEMM launcher
// ------------------------------------------------------------------------------------------------------------
struct GemmLaunch { GemmTcParams p; dim3 grid; size_t smem; int a_mn, b_mn; };

static int prepare_gemm(GemmLaunch* L, const float* A, int lda, const float* B, int ldb, float* C, int ldc, int M, int N, int K,
                        int a_mn, int b_mn, int splits, int epi, int BN) {
  memset(L, 0, sizeof *L);
  GemmTcParams& p = L->p;
  if (N % 16) return fail("gemm: N must be a multiple of 16");
  if (BN <= 0) {
    BN = N <= 256 ? N : 128;
    // These dense GEMMs are small (a few MFLOP): the time is launch + pipeline latency, so spread the output over as many SMs as there
    // are, as long as the N tile stays a legal MMA shape (g_gemm_auto_bn = 0 keeps the widest tile)
    if (g_gemm_auto_bn) {
      const long long mt = (M + 127) / 128;
      while (BN % 64 == 0 && mt * ((N + BN - 1) / BN) * std::max(1, splits) < g_num_sms) BN /= 2;
    }
  }
  if (BN > 256 || BN % 16) return fail("gemm: bad BN");
  if (b_mn && (BN % 32)) return fail("gemm: MN-major B needs BN % 32 == 0");
  if (a_mn) CIB_TRY(make_tmap_2d(&p.tmA, A, M, K, lda, 32, 32, SWZ_TMA_MN));       // A stored [K][lda], M contiguous
  else      CIB_TRY(make_tmap_2d(&p.tmA, A, K, M, lda, 32, 128, 128));      // A stored [M][lda], K contiguous
  if (b_mn) CIB_TRY(make_tmap_2d(&p.tmB, B, N, K, ldb, 32, 32, SWZ_TMA_MN));       // B stored [K][ldb], N contiguous
  else      CIB_TRY(make_tmap_2d(&p.tmB, B, K, N, ldb, 32, BN, 128));       // B stored [N][ldb], K contiguous
  p.C = C; p.ldc = ldc; p.M = M; p.N = N; p.K = K; p.BN = BN; p.epi = epi;
  const int nk_total = (K + 31) / 32;
  if (splits < 1) splits = 1;
  if (splits > nk_total) splits = nk_total;
  const int ck = (nk_total + splits - 1) / splits;
  p.k_per_split = ck * 32;
  p.splits = (nk_total + ck - 1) / ck;
  const size_t stage = GEMM_A_STAGE_BYTES + (size_t)BN * 128;
  int stages = std::min(ck, 6);
  while (stages > 1 && stages * stage + 1024 > 200 * 1024) --stages;
  p.stages = stages;
  L->smem = stages * stage + 1024;
  const int m_tiles = (M + 127) / 128, n_tiles = (N + BN - 1) / BN;
  L->grid = dim3(m_tiles * n_tiles * p.splits);
  L->a_mn = a_mn; L->b_mn = b_mn;
  return 0;
}
static int launch_gemm(const GemmLaunch& L, cudaStream_t st) {
  if (L.a_mn && L.b_mn)       { CIB_TRY(set_smem(gemm_tc_kernel<true, true>, L.smem));   CIB_CUDA(launch_k(gemm_tc_kernel<true, true>, L.grid, dim3(192), L.smem, st, L.p)); }
  else if (L.a_mn && !L.b_mn) { CIB_TRY(set_smem(gemm_tc_kernel<true, false>, L.smem));  CIB_CUDA(launch_k(gemm_tc_kernel<true, false>, L.grid, dim3(192), L.smem, st, L.p)); }
  else if (!L.a_mn && L.b_mn) { CIB_TRY(set_smem(gemm_tc_kernel<false, true>, L.smem));  CIB_CUDA(launch_k(gemm_tc_kernel<false, true>, L.grid, dim3(192), L.smem, st, L.p)); }
  else                        { CIB_TRY(set_smem(gemm_tc_kernel<false, false>, L.smem)); CIB_CUDA(launch_k(gemm_tc_kernel<false, false>, L.grid, dim3(192), L.smem, st, L.p)); }
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// conv forward / dgrad launcher (tcgen05)
// ------------------------------------------------------------------------------------------------------------
struct ConvLaunch { ConvTcParams p; dim3 grid; size_t smem; };

static bool tc_channels_ok(int c) { return c == 16 || (c >= 32 && c % 32 == 0 && c <= 256); }   // K-major operands

// forward: src = x [n][Hi][Wi][Ci], result grid (Ho, Wo); dgrad: src = dz [n][Hz][Wz][Co], result grid = input grid.
static int prepare_conv(ConvLaunch* L, bool forward, const float* src, int n, int Hs, int Ws, int Cs,
                        const float* w, int Cin_pad, int Cout_pad, int Hr, int Wr, int pad, int stride = 1, int cls = -1) {
  // stride 2: forward reads the source with traversal stride 2; dgrad is split into four unit-stride convolutions, one per parity
  // class cls = 2 a + b of the result pixel (see ConvTcParams), each over the sub-grid of that class
  memset(L, 0, sizeof *L);
  ConvTcParams& p = L->p;
  const int KK = forward ? Cin_pad : Cout_pad;
  const int NN = forward ? Cout_pad : Cin_pad;
  if (!tc_channels_ok(KK) || !tc_channels_ok(NN) || Cs != KK) return fail("conv_tc: unsupported channel counts");
  const int cb = KK >= 32 ? 32 : 16;
  CIB_TRY(make_tmap_nhwc(&p.tmA, src, n, Hs, Ws, Cs, cb, 0, forward ? stride : 1));
  p.sstride = forward ? stride : 1;
  p.ostride = 1; p.oa = p.ob = 0; p.Hfull = Hr; p.Wfull = Wr;
  const int ca = cls >= 0 ? cls >> 1 : 0, cbb = cls >= 0 ? cls & 1 : 0;
  if (!forward && stride == 2) {
    if (cls < 0) return fail("conv_tc: strided dgrad needs a parity class");
    p.ostride = 2; p.oa = ca; p.ob = cbb;
    Hr = (Hr - ca + 1) / 2; Wr = (Wr - cbb + 1) / 2;             // sub-grid of this class
  }
  if (forward) {
    if (NN % 32) return fail("conv_tc: forward needs Cout padded to a multiple of 32 (MN-major tf32 operand)");
    CIB_TRY(make_tmap_2d(&p.tmB, w, Cout_pad, 9 * (uint64_t)Cin_pad, Cout_pad, 32, KK, SWZ_TMA_MN));
  } else {
    CIB_TRY(make_tmap_2d(&p.tmB, w, Cout_pad, 9 * (uint64_t)Cin_pad, Cout_pad, cb, NN, cb * 4));
  }
  p.b_mn = forward ? 1 : 0; p.KK = KK; p.NN = NN; p.cin_pad = Cin_pad;
  p.n_img = n; p.Ho = Hr; p.Wo = Wr;
  p.TW = Wr <= 16 ? 16 : 32;
  const int TH = 128 / p.TW;
  p.tiles_w = (Wr + p.TW - 1) / p.TW; p.tiles_h = (Hr + TH - 1) / TH;
  p.num_tiles = n * p.tiles_w * p.tiles_h;
  p.ntaps = 0;
  for (int kh = 0; kh < 3; ++kh)
    for (int kw = 0; kw < 3; ++kw) {
      int t = p.ntaps;
      if (!forward && stride == 2) {
        // result pixel (2 i + a, 2 j + b) = source (i + (a + pad - kh) / 2, j + (b + pad - kw) / 2) where those are integers
        if (((ca + pad - kh) & 1) || ((cbb + pad - kw) & 1)) continue;
        p.tap_dh[t] = (ca + pad - kh) / 2; p.tap_dw[t] = (cbb + pad - kw) / 2;      // exact: the numerators are even
      } else if (forward) { p.tap_dh[t] = kh - pad; p.tap_dw[t] = kw - pad; }
      else                { p.tap_dh[t] = pad - kh; p.tap_dw[t] = pad - kw; }
      p.tap_row[t] = (kh * 3 + kw) * Cin_pad;
      ++p.ntaps;
    }
  size_t stage = (size_t)128 * KK * 4 + (size_t)KK * NN * 4;
  int stages = 8;
  while (stages > 2 && stages * stage + 1024 > 200 * 1024) --stages;
  if (stages * stage + 1024 > 220 * 1024) {
    // wide layer: one whole tap does not fit twice -> a stage is one 32-channel K chunk of a tap (see conv_tc.cuh)
    if (KK % 32 || KK < 64) return fail("conv_tc: stage does not fit in shared memory");
    p.kchunk = 1;
    stage = (size_t)128 * 128 + (size_t)32 * NN * 4;
    stages = 8;
    while (stages > 2 && stages * stage + 1024 > 200 * 1024) --stages;
    // the filter boxes shrink to 32 K rows
    if (forward) CIB_TRY(make_tmap_2d(&p.tmB, w, Cout_pad, 9 * (uint64_t)Cin_pad, Cout_pad, 32, 32, SWZ_TMA_MN));
  }
  p.stages = stages;
  L->smem = stages * stage + 1024;
  L->grid = dim3(std::min(p.num_tiles, g_num_sms));
  return 0;
}
static int launch_conv(const ConvLaunch& L, cudaStream_t st) {
  CIB_TRY(set_smem(conv_tc_kernel, L.smem));
  conv_tc_kernel<<<L.grid, 192, L.smem, st>>>(L.p);
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// conv forward / dgrad, halo-copy producer (conv_tc2.cuh).  Falls back to the TMA-per-tap kernel when the filter
// does not fit in shared memory next to two A stages.
// ------------------------------------------------------------------------------------------------------------
struct Conv2Launch { ConvTc2Params p; dim3 grid; size_t smem; bool ok; };

static bool conv2_fits(int KK, int NN, int* stages_out) {
  const size_t RB = KK >= 32 ? 128 : 64;
  const size_t stage = 3 * (size_t)C2_ROWS * RB;
  const size_t b = ((size_t)9 * KK * NN * 4 + 1023) & ~(size_t)1023;
  int stages = 4;
  while (stages > 1 && b + stages * stage + 1024 > (size_t)CIB_MAX_DYN_SMEM - 2048) --stages;
  if (b + stages * stage + 1024 > (size_t)CIB_MAX_DYN_SMEM - 2048) return false;
  if (stages_out) *stages_out = stages;
  return stages >= 2 || KK <= 32;
}

static int prepare_conv2(Conv2Launch* L, bool forward, const float* src, int n, int Hs, int Ws, int Cs,
                         const float* w, int Cin_pad, int Cout_pad, int Hr, int Wr, int pad) {
  memset(L, 0, sizeof *L);
  ConvTc2Params& p = L->p;
  const int KK = forward ? Cin_pad : Cout_pad;
  const int NN = forward ? Cout_pad : Cin_pad;
  if (!tc_channels_ok(KK) || !tc_channels_ok(NN) || Cs != KK) return fail("conv_tc2: unsupported channel counts");
  int stages = 0;
  if (!conv2_fits(KK, NN, &stages)) return fail("conv_tc2: filter does not fit in shared memory");
  if (forward) {
    if (NN % 32) return fail("conv_tc2: forward needs Cout padded to a multiple of 32");
    CIB_TRY(make_tmap_2d(&p.tmB, w, Cout_pad, 9 * (uint64_t)Cin_pad, Cout_pad, 32, KK, SWZ_TMA_MN));
  } else {
    const int cb = KK >= 32 ? 32 : 16;
    CIB_TRY(make_tmap_2d(&p.tmB, w, Cout_pad, 9 * (uint64_t)Cin_pad, Cout_pad, cb, NN, cb * 4));
  }
  p.src = src; p.Hs = Hs; p.Ws = Ws; p.b_mn = forward ? 1 : 0; p.KK = KK; p.NN = NN;
  p.n_img = n; p.Ho = Hr; p.Wo = Wr;
  p.tiles_w = (Wr + 15) / 16; p.tiles_h = (Hr + 7) / 8; p.num_tiles = n * p.tiles_w * p.tiles_h;
  const int mn = forward ? -pad : pad - 2;
  p.org_dh = mn; p.org_dw = mn;
  for (int kh = 0; kh < 3; ++kh)
    for (int kw = 0; kw < 3; ++kw) {
      const int t = kh * 3 + kw;
      p.tap_row[t] = t * Cin_pad;
      p.tap_ih[t] = forward ? kh : 2 - kh;
      p.tap_iw[t] = forward ? kw : 2 - kw;
    }
  p.stages = stages;
  const size_t RB = KK >= 32 ? 128 : 64;
  L->smem = (((size_t)9 * KK * NN * 4 + 1023) & ~(size_t)1023) + stages * 3 * (size_t)C2_ROWS * RB + 1024;
  L->grid = dim3(std::min(p.num_tiles, g_num_sms));
  L->ok = true;
  return 0;
}
static int launch_conv2(const Conv2Launch& L, cudaStream_t st) {
  CIB_CUDA(launch_k(conv_tc2_kernel, L.grid, dim3(C2_THREADS), L.smem, st, L.p));
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// One entry point for the engine and the stand-alone ops: picks the producer, fills the epilogue, launches.
struct ConvEpilogue {
  int epi = CONV_EPI_FWD;
  const float* bias = nullptr; float* out = nullptr; int pool = 0; float drop_rate = 0.f; unsigned int layer_id = 0;
  const HyperBlock* hb = nullptr;
  const float* act_prev = nullptr; float* dz_out = nullptr; int Hf = 0, Wf = 0; float scale = 1.f;
};
static int run_conv(bool forward, const float* src, int n, int Hs, int Ws, int Cs, const float* w, int Cin_pad, int Cout_pad,
                    int Hr, int Wr, int pad, const ConvEpilogue& e, cudaStream_t st, int stride = 1) {
  const int KK = forward ? Cin_pad : Cout_pad, NN = forward ? Cout_pad : Cin_pad;
  auto fill = [&](auto& P) {
    P.epi = e.epi; P.bias = e.bias; P.out = e.out; P.pool = e.pool; P.drop_rate = e.drop_rate; P.layer_id = e.layer_id;
    P.hb = e.hb; P.act_prev = e.act_prev; P.dz_out = e.dz_out; P.Hf = e.Hf; P.Wf = e.Wf; P.scale = e.scale;
  };
  if (stride == 1 && g_conv_version == 2 && conv2_fits(KK, NN, nullptr)) {
    Conv2Launch L;
    CIB_TRY(prepare_conv2(&L, forward, src, n, Hs, Ws, Cs, w, Cin_pad, Cout_pad, Hr, Wr, pad));
    fill(L.p);
    return launch_conv2(L, st);
  }
  if (stride == 2 && !forward) {
    // four parity classes, disjoint output pixels: back to back on the same stream
    if (e.epi != CONV_EPI_DGRAD_MASK) return fail("conv_tc: strided dgrad supports the mask epilogue only");
    for (int cls = 0; cls < 4; ++cls) {
      ConvLaunch L;
      CIB_TRY(prepare_conv(&L, false, src, n, Hs, Ws, Cs, w, Cin_pad, Cout_pad, Hr, Wr, pad, 2, cls));
      fill(L.p);
      if (L.p.num_tiles > 0) CIB_TRY(launch_conv(L, st));
    }
    return 0;
  }
  ConvLaunch L;
  CIB_TRY(prepare_conv(&L, forward, src, n, Hs, Ws, Cs, w, Cin_pad, Cout_pad, Hr, Wr, pad, stride));
  fill(L.p);
  return launch_conv(L, st);
}

// ------------------------------------------------------------------------------------------------------------
// conv wgrad launcher (tcgen05)
// ------------------------------------------------------------------------------------------------------------
struct WgradLaunch { WgradTcParams p; dim3 grid; size_t smem; };

static int prepare_wgrad(WgradLaunch* L, const float* x, const float* dz, float* dW, float* db, int n, int Hi, int Wi, int CI,
                         int Ho, int Wo, int CO, int pad, int stride = 1) {
  memset(L, 0, sizeof *L);
  WgradTcParams& p = L->p;
  if (CO % 32 || CO > 256) return fail("wgrad_tc: CO must be a multiple of 32 (<= 256)");
  int CIm = CI;                                   // MMA N
  if (CI == 16) {
    // 16-channel X: read rows of 32 floats = this pixel + its right neighbour (W stride stays 64 B); the MMA then
    // produces 32 columns per tap of which the first 16 are this tap.  The caller keeps >= 64 B of slack after x.
    CIm = 32;
    const uint64_t dims[4] = {32, (uint64_t)Wi, (uint64_t)Hi, (uint64_t)n};
    const uint64_t st[3] = {64, (uint64_t)Wi * 64, (uint64_t)Hi * Wi * 64};
    const uint32_t box[4] = {32, 16, 2, 1};
    CIB_TRY(make_tmap(&p.tmB, x, 4, dims, st, box, SWZ_TMA_MN));
  } else {
    if (CI % 32 || CI > 256) return fail("wgrad_tc: CI must be 16 or a multiple of 32 (<= 256)");
    CIB_TRY(make_tmap_nhwc(&p.tmB, x, n, Hi, Wi, CI, 32, SWZ_TMA_MN, stride));
  }
  if (stride != 1 && CI == 16) return fail("wgrad_tc: strided layers need CI to be a multiple of 32");
  p.xstride = stride;
  CIB_TRY(make_tmap_nhwc(&p.tmA, dz, n, Ho, Wo, CO, 32, SWZ_TMA_MN));
  p.CO = CO; p.CI = CIm; p.ci_out = CI; p.n_img = n; p.Ho = Ho; p.Wo = Wo; p.pad = pad;
  p.chunks_w = (Wo + 15) / 16; p.chunks_h = (Ho + 1) / 2; p.num_chunks = n * p.chunks_w * p.chunks_h;
  if (9 * CIm + 32 <= 512) p.taps_per_group = 9;
  else if (3 * CIm + 32 <= 512) p.taps_per_group = 3;
  else p.taps_per_group = 1;
  p.tap_groups = 9 / p.taps_per_group;
  p.m_tiles = (CO + 127) / 128;
  const int per_slice_ctas = p.tap_groups * p.m_tiles;
  p.n_slices = std::max(1, std::min(p.num_chunks, g_num_sms / per_slice_ctas));
  const size_t a_bytes = (size_t)(std::min(CO, 128) / 32) * 4096;
  const size_t b_bytes = (size_t)p.taps_per_group * (CIm / 32) * 4096;
  const size_t stage = a_bytes + std::max<size_t>(b_bytes, 16384);
  int stages = 6;
  while (stages > 2 && stages * stage + 3072 > 200 * 1024) --stages;
  if (stages * stage + 3072 > 220 * 1024) return fail("wgrad_tc: stage does not fit in shared memory");
  p.stages = stages;
  p.dW = dW; p.db = db;
  L->smem = stages * stage + 1024 + 2048;
  L->grid = dim3(p.n_slices * per_slice_ctas);
  return 0;
}
static int launch_wgrad(const WgradLaunch& L, cudaStream_t st) {
  CIB_TRY(set_smem(wgrad_tc_kernel, L.smem));
  wgrad_tc_kernel<<<L.grid, 192, L.smem, st>>>(L.p);
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// conv wgrad, halo-copy producer (wgrad_tc2.cuh) with fall-back to the TMA-per-tap kernel
// ------------------------------------------------------------------------------------------------------------
static bool wgrad2_fits(int CI, int CO, int* stages_out, size_t* smem_out) {
  if (CO % 32 || CO > 256 || !(CI == 16 || (CI % 32 == 0 && CI <= 256))) return false;
  const int cib = CI >= 32 ? CI / 32 : 1;
  if (cib & (cib - 1)) return false;                 // the loader decodes (pixel, piece) with shifts
  const int nmb = 3 * cib, ngroups = (nmb + 3) / 4;
  if ((3 * ngroups + 1) * CO > 512) return false;
  const size_t reg = (size_t)C2_ROWS * 128, stage = nmb * reg + (size_t)(CO / 32) * 128 * 128;
  int stages = 3;
  const size_t fixed = 1024 + 4096 + reg;
  while (stages > 1 && stages * stage + fixed > (size_t)CIB_MAX_DYN_SMEM - 2048) --stages;
  if (stages * stage + fixed > (size_t)CIB_MAX_DYN_SMEM - 2048) return false;
  if (stages_out) *stages_out = stages;
  if (smem_out) *smem_out = stages * stage + fixed;
  return true;
}
static int run_wgrad(const float* x, const float* dz, float* dW, float* db, int n, int Hi, int Wi, int CI, int Ho, int Wo, int CO,
                     int pad, cudaStream_t st, int stride = 1) {
  int stages = 0; size_t smem = 0;
  if (stride == 1 && g_conv_version == 2 && wgrad2_fits(CI, CO, &stages, &smem)) {
    WgradTc2Params p{};
    p.x = x; p.dz = dz; p.dW = dW; p.db = db; p.Hi = Hi; p.Wi = Wi; p.ci_mem = CI; p.n_img = n; p.Ho = Ho; p.Wo = Wo; p.CO = CO; p.pad = pad;
    p.tiles_w = (Wo + 15) / 16; p.tiles_h = (Ho + 7) / 8; p.num_tiles = n * p.tiles_w * p.tiles_h; p.stages = stages;
    const int grid = std::min(p.num_tiles, g_num_sms);
    CIB_CUDA(launch_k(wgrad_tc2_kernel, dim3(grid), dim3(W2_THREADS), smem, st, p));
    CIB_CUDA(cudaGetLastError());
    return 0;
  }
  WgradLaunch L;
  CIB_TRY(prepare_wgrad(&L, x, dz, dW, db, n, Hi, Wi, CI, Ho, Wo, CO, pad, stride));
  return launch_wgrad(L, st);
}

// ------------------------------------------------------------------------------------------------------------
// CUDA-core launchers
// ------------------------------------------------------------------------------------------------------------
static int launch_conv_direct(ConvDirectParams p, cudaStream_t st) {
  if (p.CO % 8) return fail("conv_direct: CO must be a multiple of 8");
  const size_t wbytes = (size_t)p.KH * p.KW * p.CI * p.CO * 4;
  p.w_in_smem = wbytes <= 40 * 1024 ? 1 : 0;
  const int Hs = p.pool ? p.Ho / 2 : p.Ho, Ws = p.pool ? p.Wo / 2 : p.Wo;
  const long long total = (long long)p.n_img * Hs * Ws * (p.CO / 8);
  const int blocks = (int)std::min<long long>((total + 255) / 256, (long long)g_num_sms * 8);
  conv_fwd_direct_kernel<<<blocks, 256, p.w_in_smem ? wbytes : 0, st>>>(p);
  CIB_CUDA(cudaGetLastError());
  return 0;
}
static int launch_wgrad_first(const WgradFirstParams& p, cudaStream_t st) {
  if (p.CO > 256 || 256 % p.CO) return fail("wgrad_first: CO must divide 256");
  const int groups = 256 / p.CO;
  const size_t smem = (size_t)groups * (9 * p.CI + 1) * p.CO * 4;
  const int blocks = g_num_sms * 4;
  if (p.CI == 1) { CIB_TRY(set_smem(conv_wgrad_first_kernel<1>, smem)); conv_wgrad_first_kernel<1><<<blocks, 256, smem, st>>>(p); }
  else if (p.CI == 3) { CIB_TRY(set_smem(conv_wgrad_first_kernel<3>, smem)); conv_wgrad_first_kernel<3><<<blocks, 256, smem, st>>>(p); }
  else return fail("wgrad_first: CI must be 1 or 3");
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// First conv layer, shared-memory tiled kernels: 3x3 / stride 1 / pooled / CI in {1,3} / small CO.
static bool first_tiled_ok(const CibBlock& b) {
  return b.ksize == 3 && b.stride == 1 && b.pool == 1 && (b.Ci == 1 || b.Ci == 3) && b.Co % 8 == 0 && b.Co <= 64 &&
         256 % b.Co == 0;
}
static int launch_conv_first(ConvFirstParams p, bool wgrad, cudaStream_t st) {
  p.tiles_w = (p.Wo + 15) / 16; p.tiles_h = (p.Ho + 15) / 16;
  const int ntiles = p.n_img * p.tiles_w * p.tiles_h;
  if (!wgrad) {
    const int grid = std::min(ntiles, g_num_sms * 8);
    if (p.CI == 1) CIB_CUDA(launch_k(conv_first_fwd_kernel<1>, dim3(grid), dim3(128), 0, st, p));
    else CIB_CUDA(launch_k(conv_first_fwd_kernel<3>, dim3(grid), dim3(128), 0, st, p));
  } else {
    // one 1024-thread block (four independent tile groups) per SM, see the kernel
    const int grid = std::min((ntiles + CFW_GROUPS - 1) / CFW_GROUPS, g_num_sms);
    // tile buffers of the four groups; reused at the end as [32 warps][(9 CI + 1) x CO] reduction slabs
    const size_t smem = std::max((size_t)CFW_GROUPS * (18 * 18 * p.CI + 64 * p.CO) + (size_t)(9 * p.CI + 1) * p.CO,
                                 (size_t)8 * CFW_GROUPS * (9 * p.CI + 1) * p.CO) * sizeof(float);
    if (p.CI == 1) CIB_CUDA(launch_k(conv_first_wgrad_kernel<1>, dim3(grid), dim3(256 * CFW_GROUPS), smem, st, p));
    else CIB_CUDA(launch_k(conv_first_wgrad_kernel<3>, dim3(grid), dim3(256 * CFW_GROUPS), smem, st, p));
  }
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// first layer on tensor cores (conv1_tc.cuh)
// ------------------------------------------------------------------------------------------------------------
// g_conv1_tc: 0 = only where the tiled CUDA-core kernels do not apply (first layer WITHOUT pooling, e.g. mnist.py:49: 16.5 / 25 us
// against 28 / 37 us for the direct kernels at 128 x 28x28x1 -> 32), 1 = wherever it is applicable.  For the pooled 64x64x3 -> 16
// layer the tensor-core version is correct but slower than the CUDA cores (measured 51 / 57 us against 29 / 41 us): a 128-pixel
// tile is only 4..16 tiny MMAs, and the per-tile hand-shakes, the ~40-cycle issue cost of each tcgen05.mma and the serial
// instruction latency of the producer warps dominate.
static int g_conv1_tc = 0;
static bool first_tiled_ok(const CibBlock& b);
static bool conv1_tc_ok(const CibBlock& b) {
  const bool shape_ok = b.ksize == 3 && b.stride == 1 && (b.Ci == 1 || b.Ci == 3) && (b.Co == 16 || b.Co == 32 || b.Co == 64) &&
                        (b.pad == 0 || b.pad == 1) && (!b.pool || (b.Ho % 2 == 0 && b.Wo % 2 == 0));
  return shape_ok && (g_conv1_tc || !first_tiled_ok(b));
}
static int launch_conv1_tc(Conv1TcParams p, bool wgrad, cudaStream_t st) {
  p.tiles_w = (p.Wo + 15) / 16; p.tiles_h = (p.Ho + 7) / 8; p.num_tiles = p.n_img * p.tiles_w * p.tiles_h;
  const int grid = std::min(p.num_tiles, g_num_sms);
  { static const int dbg = [] { const char* d = getenv("CIB_C1_DEBUG"); return d ? atoi(d) : 0; }(); p.debug = dbg; }   // experiments only
  if (!wgrad) {
    p.stages = 8;
    const size_t smem = 1024 + (((size_t)p.CO * 128 + 1023) & ~(size_t)1023) + (size_t)p.stages * C1_TILE_BYTES;
    CIB_CUDA(launch_k(conv1_tc_fwd_kernel, dim3(grid), dim3(C1_FWD_THREADS), smem, st, p));
  } else {
    const size_t stage = (size_t)C1_TILE_BYTES + (size_t)p.CO * 512;          // A^T tile + dZ^T tile
    p.stages = (int)std::min<size_t>(8, ((size_t)190 * 1024 - C1_TILE_BYTES) / stage);
    const size_t smem = 1024 + (size_t)p.stages * stage + C1_TILE_BYTES;      // slack: rows 32..127 of the last A^T block
    conv1_tc_wgrad_kernel<<<grid, C1_WG_THREADS, smem, st>>>(p);
  }
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// Kernels that are meant to share an SM with a convolution CTA (which runs with the maximum shared-memory carve-out) must ask for
// the same carve-out: an SM cannot hold CTAs of two kernels with different L1/shared splits at the same time, so with the default
// (small shared, large L1) the all-reduce CTAs waited ~25 us for an SM to drain instead of slipping in next to the convolution.
template <typename K>
static int prefer_max_shared(K kernel) {
  CIB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  return 0;
}

static int init_kernel_attributes() {
  CIB_TRY(prefer_max_shared(push_allreduce_opt_kernel)); CIB_TRY(prefer_max_shared(ll_allreduce_opt_kernel));
  CIB_TRY(prefer_max_shared(allreduce_opt_kernel)); CIB_TRY(prefer_max_shared(stamp_kernel));
  CIB_TRY(prefer_max_shared(optimizer_kernel));
  CIB_TRY(raise_smem(conv1_tc_fwd_kernel)); CIB_TRY(raise_smem(conv1_tc_wgrad_kernel));
  CIB_TRY(raise_smem(gemm_tc_kernel<true, true>)); CIB_TRY(raise_smem(gemm_tc_kernel<true, false>));
  CIB_TRY(raise_smem(gemm_tc_kernel<false, true>)); CIB_TRY(raise_smem(gemm_tc_kernel<false, false>));
  CIB_TRY(raise_smem(conv_tc_kernel)); CIB_TRY(raise_smem(wgrad_tc_kernel)); CIB_TRY(raise_smem(conv_tc2_kernel)); CIB_TRY(raise_smem(wgrad_tc2_kernel));
  CIB_TRY(raise_smem(conv_wgrad_first_kernel<1>)); CIB_TRY(raise_smem(conv_wgrad_first_kernel<3>));
  CIB_TRY(raise_smem(conv_first_wgrad_kernel<1>)); CIB_TRY(raise_smem(conv_first_wgrad_kernel<3>));
  CIB_TRY(raise_smem(conv_fwd_direct_kernel)); CIB_TRY(raise_smem(head_kernel<1>)); CIB_TRY(raise_smem(head_kernel<16>));
  CIB_TRY(raise_smem(head_mc_kernel<1>)); CIB_TRY(raise_smem(head_mc_kernel<16>));
  return 0;
}

}  // namespace cib
