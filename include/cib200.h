/* C ABI of libcib200.so -- the sm_100a runtime of cori_intml_examples_b200.
 *
 * Everything here is callable through ctypes (that is how the Python package binds it) or from C/C++.
 * All functions return 0 on success and a non-zero code on failure; cib_last_error() describes the failure.
 * Device pointers are plain `void*` / `float*` values as seen by the calling process.
 */
#ifndef CIB200_H_
#define CIB200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- runtime ---------------------------------- */
int cib_init(int device);                       /* select device, create the error mailbox (idempotent)        */
const char* cib_last_error(void);
int cib_device_error(unsigned int out[4]);      /* code written by a kernel that trapped on a bounded wait     */
int cib_version(void);
/* Process-wide exclusive section: take it around device-wide synchronisation or device (de)allocation done outside this
 * library when other threads of the process may be capturing CUDA graphs (several trials per engine process). */
int cib_exclusive_begin(void);
int cib_exclusive_end(void);

/* ---------------------------------------------------------------- model program ---------------------------- */
/* One fused block of the layer chain.  Built by the Python planner (ops/plan.py).
 * Reference layer stacks: mnist.py:48-56, rpv.py:42-59. */
typedef struct CibBlock {
  int kind;                /* 0 = conv(+pool)(+dropout), 1 = hidden dense(+dropout), 2 = output head            */
  int path;                /* 0 = CUDA-core kernels, 1 = tcgen05 kernels                                        */
  int Hi, Wi, Ci;          /* conv: input grid and channels as stored in memory                                 */
  int Ho, Wo, Co;          /* conv: output grid before pooling, padded output channels                          */
  int ksize, stride, pad;  /* conv: 3 / 1 / 0|1 on the tcgen05 path                                             */
  int pool;                /* conv: 1 = fused 2x2 max-pool                                                      */
  int K, N;                /* dense/head: input features in memory, (padded) output features                    */
  int head_kind;           /* head: 0 = sigmoid + binary CE, 1 = softmax + categorical CE                       */
  float drop_rate;         /* dropout applied to this block's output (0 = none)                                 */
  long long w_off, b_off;  /* offsets (floats) of kernel and bias in the flat parameter buffer                  */
} CibBlock;

typedef struct CibHyper {
  float lr, beta1, beta2, eps, rho, momentum, schedule_decay, decay;
  float m_schedule;
  unsigned int iterations;
  unsigned int seed;
} CibHyper;

typedef struct CibMetrics {
  float loss_sum, correct_sum, sample_sum, step_loss, step_acc;
  unsigned int iterations;
  float m_schedule;
  unsigned int batch_cursor;
} CibMetrics;

typedef void* cib_engine_t;
typedef void* cib_comm_t;

/* opt_kind: 0 SGD, 1 Adam, 2 Nadam, 3 Adadelta (Keras 2.2 update rules). */
int cib_engine_create(cib_engine_t* out, const CibBlock* blocks, int nblocks, int batch,
                      int in_h, int in_w, int in_c, long long nparam_padded, int opt_kind, int flags);
int cib_engine_destroy(cib_engine_t e);
/* which: 0 params, 1 grads, 2 slot0, 3 slot1 (flat padded buffers, host <-> device) */
int cib_engine_upload(cib_engine_t e, int which, const float* host, long long n);
int cib_engine_download(cib_engine_t e, int which, float* host, long long n);
int cib_engine_set_hyper(cib_engine_t e, const CibHyper* h);
int cib_engine_set_lr(cib_engine_t e, float lr);
int cib_engine_get_metrics(cib_engine_t e, CibMetrics* out);      /* synchronises the engine stream          */
int cib_engine_reset_metrics(cib_engine_t e);
/* Device-resident dataset: x [n][H][W][C] fp32, y [n][label_w] fp32.  perm (host int32, may be NULL = identity)
 * is copied to the device; the cursor restarts at 0.  Reference fit(): rpv.py:103-106. */
int cib_engine_set_dataset(cib_engine_t e, const float* dev_x, const float* dev_y, int n, int label_w);
int cib_engine_set_perm(cib_engine_t e, const int* host_perm, int n);
int cib_engine_train_steps(cib_engine_t e, int nsteps);           /* graph replay, no host sync               */
int cib_engine_eval_steps(cib_engine_t e, int nsteps, float* dev_probs_out);
/* nsteps train steps bracketed by CUDA events on the engine stream; *ms_out = device time in milliseconds. */
int cib_engine_time_steps(cib_engine_t e, int nsteps, float* ms_out);
/* The same, with `preroll` untimed steps issued immediately before the first event (no host synchronisation in between): under
 * data parallelism their collectives align the ranks on the device, so short timed regions do not measure process start skew. */
int cib_engine_time_steps_aligned(cib_engine_t e, int preroll, int nsteps, float* ms_out);
/* 1: the batch cursor wraps to 0 instead of running past the bound dataset (timing loops longer than one pass); 0 = default. */
int cib_engine_set_wrap(cib_engine_t e, int on);
/* End-to-end step on a host batch (pinned memory): H2D copy, train step, D2H of the batch loss.
 * The loss of this step is valid after cib_engine_sync or once the NEXT call returns (pipelined). */
int cib_engine_train_batch_host(cib_engine_t e, const float* host_x, const float* host_y, int n_valid,
                                float* host_loss_out);
/* Zero-copy result of host-fed steps: pinned host ring of {loss, accuracy} pairs, slot = (optimizer step before the step) & (entries-1),
 * written by the loss head itself (pass host_loss_out = NULL to cib_engine_train_batch_host); valid after cib_engine_sync. */
int cib_engine_loss_ring(cib_engine_t e, float** host_ring, int* entries);
int cib_engine_sync(cib_engine_t e);
int cib_engine_launch_count(cib_engine_t e, int* train_step_kernels, int* eval_step_kernels);
/* Debug build of the buffers (environment CIB_DEBUG_BOUNDS=1 before cib_init): 256-byte canary bands around every engine buffer.
 * *bad = overwritten canary bytes (0 = intact, -1 = canaries not enabled); cib_last_error() names the first damaged buffer. */
int cib_engine_check_canaries(cib_engine_t e, long long* bad);
/* Raw device pointers of the engine's buffers (for tests): which 0..3 as above, 4 = batch x, 5 = batch y,
 * 6 = probs, 7 = sample weights, 100+i = activation of block i, 200+i = dz of block i. */
int cib_engine_buffer(cib_engine_t e, int which, void** ptr, long long* nfloats);
/* Run forward(+backward) once on whatever is in the batch buffers, without the optimizer (for tests). */
int cib_engine_forward_backward(cib_engine_t e, int training, int backward);
/* Trials packed on one GPU: high = 1 re-creates the engine's compute streams at the greatest CUDA stream priority (0: the least)
 * and re-captures its graphs at the next step; get returns the priority the main stream really has and the device's range. */
int cib_engine_set_stream_priority(cib_engine_t e, int high);
int cib_engine_get_stream_priority(cib_engine_t e, int* prio, int* least, int* greatest);

/* ---------------------------------------------------------------- collectives ------------------------------ */
/* Same-process communicator over `world` devices (peer access) -- tests and single-process multi-GPU. */
int cib_comm_create_local(cib_comm_t* out, cib_engine_t* engines, int world);
/* Multi-process: export this engine's arena, then attach the peers' handles (64 bytes each, rank order). */
int cib_engine_ipc_handle(cib_engine_t e, unsigned char handle[64]);
int cib_comm_create_ipc(cib_comm_t* out, cib_engine_t e, int rank, int world, const unsigned char* handles);
int cib_comm_destroy(cib_comm_t c);
/* After this, cib_engine_train_steps runs the fused all-reduce + optimizer kernel instead of the local optimizer.
 * mode: 0 auto, 1 one-shot, 2 two-shot. */
int cib_engine_attach_comm(cib_engine_t e, cib_comm_t c, int rank_in_comm, int mode);
int cib_engine_broadcast_state(cib_engine_t e, int root);         /* params + optimizer slots from root       */
/* Two-shot updates shard the optimizer moments (rank r updates slice r only).  Collective: every rank pulls the other slices
 * from their owners, after which cib_engine_download(2|3) is complete on every rank (checkpoints, engine re-creation). */
int cib_engine_gather_slots(cib_engine_t e);
/* NVLS on the multi-process training path: the engine's gradient buffer is bound into one multicast object shared by all ranks, and
 * large buckets are reduced inside the NVSwitch (multimem.ld_reduce) by the fused all-reduce + optimizer kernel.  One collective
 * set-up per engine, driven by the caller with a host barrier between the steps: 0 size (all) / 1 rank 0 creates and exports a POSIX
 * fd through fd_out, the others import fd_in / 2 add the own device (all) / 3 bind + map (all) / 9 undo.  A failing step (no multicast
 * on the platform) leaves the peer-to-peer path in place. */
int cib_engine_nvls_step(cib_engine_t e, int step, int world, int fd_in, int* fd_out);
int cib_engine_nvls_active(cib_engine_t e);
/* globaltimer stamps of the fused push all-reduce of the last step (two buckets x 8 slots) and of 5 step-level markers
 * (slots 16..20); out32 receives 32 values; see csrc/comm_api.cuh. */
int cib_engine_ar_trace(cib_engine_t e, int enable, unsigned long long* out32);
/* A/B switches and limits of the gradient all-reduce: pull kernel instead of the push kernel; CTAs of the push kernel (<= 64);
 * time-out of every cross-GPU wait (default 180 s, also CIB_PEER_TIMEOUT_S). */
int cib_set_ar_pull(int on);
/* 1 (default): kernels of the step's main stream use programmatic dependent launch (their prologue overlaps the previous kernel's
 * tail); 0: ordinary stream order.  Takes effect at the next graph capture.  Also CIB_PDL. */
int cib_set_pdl(int on);
int cib_set_ar_ll(int on);          /* 1 (default): small one-shot buckets use the flag-in-data (LL) protocol */
int cib_set_par_ctas(int n);
int cib_set_peer_timeout_s(double seconds);

/* Stand-alone symmetric buffers for the all-reduce bandwidth sweep and the metric average. */
typedef void* cib_symbuf_t;
int cib_symbuf_create(cib_symbuf_t* out, int device, long long nfloats);
int cib_symbuf_ipc_handle(cib_symbuf_t s, unsigned char handle[64]);
int cib_symbuf_attach_ipc(cib_symbuf_t s, int rank, int world, const unsigned char* handles);
int cib_symbuf_attach_local(cib_symbuf_t* bufs, int world);
int cib_symbuf_ptr(cib_symbuf_t s, void** dev_ptr);
int cib_symbuf_allreduce(cib_symbuf_t s, long long nfloats, int average, int mode, void* stream);
int cib_symbuf_broadcast(cib_symbuf_t s, long long nfloats, int root, void* stream);
int cib_symbuf_destroy(cib_symbuf_t s);

/* NVLS all-reduce: reduction and broadcast inside the NVSwitch (multimem.ld_reduce / multimem.st) on a multicast-mapped
 * buffer shared by GPUs 0..world-1 of THIS process.  cib_nvls_create_local fails with a message when the platform does not
 * expose multicast; use the cib_symbuf_* peer-to-peer collectives then.  cib_nvls_ptr gives rank's own (unicast) view of
 * the buffer; call cib_nvls_allreduce once per rank (each on its own device / stream). */
typedef void* cib_nvls_t;
int cib_nvls_supported(int device, int* supported);
int cib_nvls_create_local(cib_nvls_t* out, int world, long long nfloats);
int cib_nvls_ptr(cib_nvls_t s, int rank, void** dev_ptr);
int cib_nvls_allreduce(cib_nvls_t s, int rank, long long nfloats, int average, void* stream);
int cib_nvls_destroy(cib_nvls_t s);

/* ---------------------------------------------------------------- stand-alone ops (tests, benches) --------- */
/* C[M,N] (+)= A * B on tcgen05.  a_mn / b_mn select MN-major operands (see csrc/gemm_tc.cuh).
 * epi: 0 store, 1 red.add.  All pointers are device pointers; K, N multiples of 32 for MN-major operands. */
int cib_gemm_tf32(const float* A, const float* B, float* C, int M, int N, int K, int a_mn, int b_mn,
                  int splits, int epi, void* stream);
/* Hidden Dense layer forward as the engine runs it (reference rpv.py:55 Dense(relu)): out[M][N] = relu(A[M][K] * W[K][N] + bias),
 * split-K tcgen05 GEMM accumulating into the fp32 scratch z_scratch [M][N], which must be zero on entry and is zero again on
 * return.  N a multiple of 32 (<= 256), K a multiple of 16. */
int cib_dense_relu_fwd(const float* A, const float* W, const float* bias, float* out, float* z_scratch, int M, int N, int K,
                       void* stream);
/* The same dense-layer forward for callers that cannot hand over a scratch buffer: HPO trials sharing a GPU (a
 * cudaMalloc/cudaFree would synchronise the device and stall the neighbouring trials' graph replays) and trials living
 * inside a fixed HBM slice given out by the scheduler.  This entry point has no counterpart in the reference.
 * The scratch is a 256 KiB array compiled into the library (held by every GPU that loads it, until exit), hence the limit
 * M * N <= 65536; larger problems are refused with an error.  The first call on a device performs a one-time set-up for
 * that device -- make it outside of any stream capture.  All calls share the one array: calls on the same stream are
 * ordered and safe, calls that may overlap on the GPU race on it.  Nothing here calls cudaMalloc or cudaFree. */
int cib_dense_relu_fwd_noalloc(const float* A, const float* W, const float* bias, float* out, int M, int N, int K,
                               void* stream);
/* 3x3 stride-1 conv forward on tcgen05: x [n][H][W][Ci] (Ci = 16 or multiple of 32), w [9][Ci][Co], fused
 * bias + ReLU (+ 2x2 max-pool).  out [n][Hs][Ws][Co]. */
int cib_conv3x3_fwd_tf32(const float* x, const float* w, const float* bias, float* out, int n, int H, int W,
                         int Ci, int Co, int pad_same, int pool, void* stream);
/* dgrad: dz [n][Ho][Wo][Co] -> masked gradient on the input grid; mode 1 = un-pool scatter into
 * dx_out [n][2*Hi'][...], mode 2 = masked store at the same resolution.  act_prev is the stored activation. */
int cib_conv3x3_dgrad_tf32(const float* dz, const float* w, const float* act_prev, float* dx_out, int n,
                           int Hi, int Wi, int Ci, int Co, int pad_same, int mode, int Hf, int Wf, float scale,
                           void* stream);
/* wgrad: dW [9][Ci][Co] += X^T dZ, db [Co] += sum dZ (both must be zero-initialised by the caller). */
int cib_conv3x3_wgrad_tf32(const float* x, const float* dz, float* dW, float* db, int n, int Hi, int Wi, int Ci,
                           int Co, int pad_same, void* stream);
/* Debug timeline: kernels stamp %globaltimer at role milestones for CTA 0 and 1 (32 slots each, ns; 0 = not hit). */
int cib_ktrace_enable(int on);
int cib_ktrace_read(unsigned long long* out64);
/* Producer selection for the conv kernels: 2 (default) = halo-copy loader warps + resident filter, 1 = TMA per tap. */
int cib_set_conv_version(int v);
/* Two buckets per step (the dense stack's gradients are all-reduced and its parameters updated while the convolution backward
 * still runs): 1 (default) under data parallelism only, 2 always, 0 never.  Takes effect at the next graph capture. */
int cib_set_early_update(int on);
/* 2 (default): loss head spread over several CTAs (partial sums added with red.add); 1: single-CTA head with a fixed summation
 * order.  Takes effect at the next graph capture. */
int cib_set_head_version(int v);
/* CUDA-core direct conv forward (any Ci; Co multiple of 8). */
/* First conv layer (Ci = 1 or 3; Co = 16, 32, 64) on tcgen05 through an im2col producer: forward with fused bias + ReLU
 * (+ 2x2 max-pool, arg-max in the 2 mantissa LSBs), and its weight/bias gradient ACCUMULATED into dW [9*Ci][Co] / db [Co]
 * from the gradient at the stored grid (pooled grid if pool) and the stored activation. */
int cib_set_conv1_tc(int on);
int cib_conv1_fwd_tc(const float* x, const float* w, const float* bias, float* out, int n, int Hi, int Wi, int Ci, int Co, int pad,
                     int pool, void* stream);
int cib_conv1_wgrad_tc(const float* x, const float* g, const float* act, float* dW, float* db, int n, int Hi, int Wi, int Ci, int Co,
                       int pad, int pool, void* stream);
int cib_conv_fwd_direct(const float* x, const float* w, const float* bias, float* out, int n, int Hi, int Wi,
                        int Ci, int Co, int ksize, int stride, int pad, int pool, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CIB200_H_ */
