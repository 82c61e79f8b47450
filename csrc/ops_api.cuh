// Stand-alone entry points for single kernels (numerics tests, micro-benchmarks, profiling).
#pragma once
#include "engine_graph.cuh"

using namespace cib;

extern "C" {

int cib_gemm_tf32(const float* A, const float* B, float* C, int M, int N, int K, int a_mn, int b_mn, int splits, int epi,
                  void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  GemmLaunch L;
  const int lda = a_mn ? M : K, ldb = b_mn ? N : K;
  CIB_TRY(prepare_gemm(&L, A, lda, B, ldb, C, N, M, N, K, a_mn, b_mn, splits, epi, 0));
  return launch_gemm(L, (cudaStream_t)stream);
}

int cib_conv3x3_fwd_tf32(const float* x, const float* w, const float* bias, float* out, int n, int H, int W, int Ci, int Co,
                         int pad_same, int pool, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  const int pad = pad_same ? 1 : 0;
  const int Ho = pad_same ? H : H - 2, Wo = pad_same ? W : W - 2;
  ConvLaunch L;
  CIB_TRY(prepare_conv(&L, true, x, n, H, W, Ci, w, Ci, Co, Ho, Wo, pad));
  L.p.epi = CONV_EPI_FWD; L.p.bias = bias; L.p.out = out; L.p.pool = pool; L.p.drop_rate = 0.f; L.p.hb = nullptr;
  return launch_conv(L, (cudaStream_t)stream);
}

int cib_conv3x3_dgrad_tf32(const float* dz, const float* w, const float* act_prev, float* dx_out, int n, int Hi, int Wi, int Ci,
                           int Co, int pad_same, int mode, int Hf, int Wf, float scale, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  const int pad = pad_same ? 1 : 0;
  const int Ho = pad_same ? Hi : Hi - 2, Wo = pad_same ? Wi : Wi - 2;
  ConvLaunch L;
  CIB_TRY(prepare_conv(&L, false, dz, n, Ho, Wo, Co, w, Ci, Co, Hi, Wi, pad));
  L.p.epi = mode; L.p.act_prev = act_prev; L.p.dz_out = dx_out; L.p.Hf = Hf; L.p.Wf = Wf; L.p.scale = scale; L.p.hb = nullptr;
  return launch_conv(L, (cudaStream_t)stream);
}

int cib_conv3x3_wgrad_tf32(const float* x, const float* dz, float* dW, float* db, int n, int Hi, int Wi, int Ci, int Co,
                           int pad_same, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  const int pad = pad_same ? 1 : 0;
  const int Ho = pad_same ? Hi : Hi - 2, Wo = pad_same ? Wi : Wi - 2;
  WgradLaunch L;
  CIB_TRY(prepare_wgrad(&L, x, dz, dW, db, n, Hi, Wi, Ci, Ho, Wo, Co, pad));
  return launch_wgrad(L, (cudaStream_t)stream);
}

int cib_conv_fwd_direct(const float* x, const float* w, const float* bias, float* out, int n, int Hi, int Wi, int Ci, int Co,
                        int ksize, int stride, int pad, int pool, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  ConvDirectParams p{};
  p.x = x; p.w = w; p.bias = bias; p.out = out; p.n_img = n; p.Hi = Hi; p.Wi = Wi; p.CI = Ci; p.CO = Co;
  p.Ho = (Hi + 2 * pad - ksize) / stride + 1; p.Wo = (Wi + 2 * pad - ksize) / stride + 1;
  p.KH = p.KW = ksize; p.stride = stride; p.pad = pad; p.pool = pool; p.drop_rate = 0.f; p.hb = nullptr;
  return launch_conv_direct(p, (cudaStream_t)stream);
}

}  // extern "C"
