// Stand-alone entry points for single kernels (numerics tests, micro-benchmarks, profiling): the ops behind
// Conv2D / Dense of mnist.py:49-56 and rpv.py:46-59 one at a time, each checked against a PyTorch fp32 reference.
#pragma once
#include "engine_graph.cuh"

using namespace cib;

extern "C" {

// In-kernel timeline (debug): enable -> kernels of CTA 0/1 stamp globaltimer into a 64-slot buffer; read copies it out.
static unsigned long long* g_ktrace_buf = nullptr;
int cib_ktrace_enable(int on) {
  if (on && !g_ktrace_buf) CIB_CUDA(cudaMalloc(&g_ktrace_buf, 64 * 8));
  if (g_ktrace_buf) CIB_CUDA(cudaMemset(g_ktrace_buf, 0, 64 * 8));
  unsigned long long* v = on ? g_ktrace_buf : nullptr;
  CIB_CUDA(cudaMemcpyToSymbol(g_ktrace, &v, sizeof v));
  return 0;
}
int cib_ktrace_read(unsigned long long* out64) {
  if (!g_ktrace_buf) return fail("ktrace not enabled");
  CIB_CUDA(cudaDeviceSynchronize());
  CIB_CUDA(cudaMemcpy(out64, g_ktrace_buf, 64 * 8, cudaMemcpyDeviceToHost));
  CIB_CUDA(cudaMemset(g_ktrace_buf, 0, 64 * 8));
  return 0;
}

// 2 = halo-copy producer when the filter fits in shared memory (default), 1 = always the TMA-per-tap kernels.
int cib_set_conv_version(int v) { g_conv_version = v; return 0; }
int cib_set_early_update(int on) { g_early_update = on; return 0; }
int cib_set_head_version(int v) { g_head_version = v; return 0; }
int cib_set_ar_pull(int on) { g_ar_pull = on; return 0; }
int cib_set_pdl(int on) { g_use_pdl = on; return 0; }
int cib_set_ar_ll(int on) { g_ar_ll = on; return 0; }
int cib_set_par_ctas(int n) { g_par_ctas = std::max(1, std::min(n, CIB_PAR_MAX_CTAS)); return 0; }
int cib_set_peer_timeout_s(double seconds) { g_peer_timeout_ns = (unsigned long long)(std::max(0.001, seconds) * 1e9); return 0; }

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
  ConvEpilogue e;
  e.epi = CONV_EPI_FWD; e.bias = bias; e.out = out; e.pool = pool;
  return run_conv(true, x, n, H, W, Ci, w, Ci, Co, Ho, Wo, pad, e, (cudaStream_t)stream);
}

int cib_conv3x3_dgrad_tf32(const float* dz, const float* w, const float* act_prev, float* dx_out, int n, int Hi, int Wi, int Ci,
                           int Co, int pad_same, int mode, int Hf, int Wf, float scale, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  const int pad = pad_same ? 1 : 0;
  const int Ho = pad_same ? Hi : Hi - 2, Wo = pad_same ? Wi : Wi - 2;
  ConvEpilogue e;
  e.epi = mode; e.act_prev = act_prev; e.dz_out = dx_out; e.Hf = Hf; e.Wf = Wf; e.scale = scale;
  return run_conv(false, dz, n, Ho, Wo, Co, w, Ci, Co, Hi, Wi, pad, e, (cudaStream_t)stream);
}

int cib_conv3x3_wgrad_tf32(const float* x, const float* dz, float* dW, float* db, int n, int Hi, int Wi, int Ci, int Co,
                           int pad_same, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  const int pad = pad_same ? 1 : 0;
  const int Ho = pad_same ? Hi : Hi - 2, Wo = pad_same ? Wi : Wi - 2;
  return run_wgrad(x, dz, dW, db, n, Hi, Wi, Ci, Ho, Wo, Co, pad, (cudaStream_t)stream);
}

// Hidden Dense layer forward exactly as the engine runs it: split-K tcgen05 GEMM accumulating with red.add into the
// fp32 scratch `z` [M][N] (must be zero on entry; it is zero again on return), then bias + ReLU -> out [M][N].
// A [M][K] row-major, W [K][N] (Keras (in, out) layout), N a multiple of 32 (<= 256), K a multiple of 16.
static int dense_relu_fwd_impl(const float* A, const float* W, const float* bias, float* out, float* z, int M, int N, int K,
                               cudaStream_t st) {
  if (N % 32 || N > 256 || K % 16 || M < 1) return fail("dense_relu_fwd: need N % 32 == 0, N <= 256, K % 16 == 0");
  GemmLaunch L;
  const int nk = (K + 31) / 32;
  const int splits = std::max(1, std::min(nk / 4, g_num_sms));
  CIB_TRY(prepare_gemm(&L, A, K, W, N, z, N, M, N, K, 0, 1, splits, GEMM_EPI_REDADD, 0));
  CIB_TRY(launch_gemm(L, st));
  bias_relu_dropout_kernel<<<grid_for((long long)M * N), 256, 0, st>>>(z, bias, out, M, N, 0.f, 0u, nullptr);
  CIB_CUDA(cudaGetLastError());
  return 0;
}
int cib_dense_relu_fwd(const float* A, const float* W, const float* bias, float* out, float* z_scratch, int M, int N, int K,
                       void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  return dense_relu_fwd_impl(A, W, bias, out, z_scratch, M, N, K, (cudaStream_t)stream);
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

// First-layer convolution on the tensor cores (conv1_tc.cuh).  x [n][Hi][Wi][Ci] (Ci = 1 or 3), w [9*Ci][Co] (Co = 16, 32, 64),
// fused bias + ReLU (+ 2x2 max-pool with arg-max bits in the two mantissa LSBs of the output).
int cib_set_conv1_tc(int on) { g_conv1_tc = on; return 0; }
int cib_conv1_fwd_tc(const float* x, const float* w, const float* bias, float* out, int n, int Hi, int Wi, int Ci, int Co, int pad,
                     int pool, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  Conv1TcParams p{};
  p.x = x; p.w = w; p.bias = bias; p.out = out; p.n_img = n; p.Hi = Hi; p.Wi = Wi; p.CI = Ci; p.CO = Co; p.pad = pad; p.pool = pool;
  p.Ho = Hi + 2 * pad - 2; p.Wo = Wi + 2 * pad - 2;
  if (!(Ci == 1 || Ci == 3) || !(Co == 16 || Co == 32 || Co == 64)) return fail("conv1_tc: Ci must be 1 or 3 and Co 16, 32 or 64");
  return launch_conv1_tc(p, false, (cudaStream_t)stream);
}
// Weight + bias gradient of that layer, ACCUMULATED into dW [9*Ci][Co] / db [Co].  g: gradient at the stored grid (pooled grid
// when pool = 1, already masked), act: the layer's stored output (supplies the arg-max bits when pool = 1).
int cib_conv1_wgrad_tc(const float* x, const float* g, const float* act, float* dW, float* db, int n, int Hi, int Wi, int Ci, int Co,
                       int pad, int pool, void* stream) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  Conv1TcParams p{};
  p.x = x; p.g = g; p.act = act; p.dW = dW; p.db = db; p.n_img = n; p.Hi = Hi; p.Wi = Wi; p.CI = Ci; p.CO = Co; p.pad = pad; p.pool = pool;
  p.Ho = Hi + 2 * pad - 2; p.Wo = Wi + 2 * pad - 2;
  if (!(Ci == 1 || Ci == 3) || !(Co == 16 || Co == 32 || Co == 64)) return fail("conv1_tc: Ci must be 1 or 3 and Co 16, 32 or 64");
  return launch_conv1_tc(p, true, (cudaStream_t)stream);
}

}  // extern "C"
