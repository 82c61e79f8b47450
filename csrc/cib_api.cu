// extern "C" surface of libcib200.so (see include/cib200.h).  Single translation unit: every kernel is compiled
// here for sm_100a.  The engine entry points implement what the reference does through model.fit / evaluate / predict
// (rpv.py:103-106, train_rpv.py:83-85, DistTrain_rpv.ipynb:407-408) on device-resident data with CUDA-graph replay.
#include <cstddef>
#include "engine_graph.cuh"

using namespace cib;

static inline Engine* eng_dev(void* e) { Engine* p = reinterpret_cast<Engine*>(e); cudaSetDevice(p->device); return p; }
#define ENG(e) (eng_dev(e))

// Data-parallel engines: a cross-GPU wait that timed out (dead or stuck peer) left a code in the device error mailbox.
static int check_peer_error(Engine* e) {
  if (!e->comm || e->comm->world < 2) return 0;
  unsigned int err[4] = {0, 0, 0, 0};
  CIB_CUDA(cudaMemcpyFromSymbol(err, g_dev_error, 16));
  if (err[0] != 0) {
    char buf[200];
    snprintf(buf, sizeof buf, "data-parallel collective timed out waiting for a peer (code 0x%08x, CTA %u, epoch %u, saw %u): a rank died or is stuck",
             err[0], err[1], err[2], err[3]);
    return fail(buf);
  }
  return 0;
}

extern "C" {

int cib_version(void) { return 100; }

const char* cib_last_error(void) { return g_last_error.c_str(); }

int cib_init(int device) {
  CIB_CUDA(cudaSetDevice(device));
  g_device = device;
  cudaDeviceProp prop;
  CIB_CUDA(cudaGetDeviceProperties(&prop, device));
  g_num_sms = prop.multiProcessorCount;
  if (prop.major != 10) return fail(std::string("cib200 needs an sm_100 GPU, found sm_") + std::to_string(prop.major * 10 + prop.minor));
  if (const char* ev = getenv("CIB_EARLY_UPDATE")) g_early_update = atoi(ev);     // A/B switches for measurements
  if (const char* ev = getenv("CIB_AR_PULL")) g_ar_pull = atoi(ev);
  if (const char* ev = getenv("CIB_PDL")) g_use_pdl = atoi(ev);
  if (const char* ev = getenv("CIB_OPT_GRID_MULT")) g_opt_grid_mult = std::max(1, atoi(ev));
  if (const char* ev = getenv("CIB_COPY_CHUNKS")) g_copy_chunks = std::max(1, std::min(8, atoi(ev)));
  if (const char* ev = getenv("CIB_DEBUG_BOUNDS")) g_debug_bounds = atoi(ev);
  if (const char* ev = getenv("CIB_DGRAD_BN")) g_dgrad_bn = atoi(ev);
  if (const char* ev = getenv("CIB_HEAD_SPC")) g_head_spc = std::max(2, atoi(ev));
  if (const char* ev = getenv("CIB_GEMM_AUTOBN")) g_gemm_auto_bn = atoi(ev);
  if (const char* ev = getenv("CIB_SERIAL_BWD")) g_serial_bwd = atoi(ev);
  if (const char* ev = getenv("CIB_FUSE_DENSE_ACT")) g_fuse_dense_act = atoi(ev);
  if (const char* ev = getenv("CIB_AR_LL")) g_ar_ll = atoi(ev);
  if (const char* ev = getenv("CIB_PAR_CTAS")) g_par_ctas = std::max(1, std::min(atoi(ev), CIB_PAR_MAX_CTAS));
  if (const char* ev = getenv("CIB_PEER_TIMEOUT_S")) g_peer_timeout_ns = (unsigned long long)(std::max(0.001, atof(ev)) * 1e9);
  CIB_TRY(ensure_driver());
  CIB_TRY(init_kernel_attributes());      // per device: function attributes live in the device's context
  return 0;
}

int cib_device_error(unsigned int out[4]) {
  // best effort: after a trap the context may be gone, in which case the copy itself fails
  cudaError_t e = cudaMemcpyFromSymbol(out, g_dev_error, 16);
  if (e != cudaSuccess) { out[0] = out[1] = out[2] = out[3] = 0xFFFFFFFFu; return fail(cudaGetErrorString(e)); }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
int cib_engine_create(cib_engine_t* out, const CibBlock* blocks, int nblocks, int batch, int in_h, int in_w, int in_c,
                      long long nparam_padded, int opt_kind, int flags) {
  if (g_device < 0) CIB_TRY(cib_init(0));
  std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
  Engine* e = new Engine();
  int rc = e->create(blocks, nblocks, batch, in_h, in_w, in_c, nparam_padded, opt_kind, flags);
  if (rc) { e->destroy(); delete e; return rc; }
  *out = e;
  return 0;
}
int cib_engine_destroy(cib_engine_t e) {
  std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
  if (e) { ENG(e)->destroy(); delete ENG(e); }
  return 0;
}
// Exclusive section for callers that are about to synchronise the whole device or (de)allocate device memory while other
// threads of the process may be capturing CUDA graphs (see g_exclusive_mu).  Recursive; always pair begin with end.
int cib_exclusive_begin(void) { g_exclusive_mu.lock(); return 0; }
int cib_exclusive_end(void) { g_exclusive_mu.unlock(); return 0; }

static float* which_buf(Engine* e, int which) {
  switch (which) { case 0: return e->params; case 1: return e->grads; case 2: return e->slot0; case 3: return e->slot1; }
  return nullptr;
}
int cib_engine_upload(cib_engine_t e_, int which, const float* host, long long n) {
  Engine* e = ENG(e_); float* d = which_buf(e, which);
  if (!d || n > e->nparam) return fail("upload: bad buffer or size");
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  CIB_CUDA(cudaMemcpy(d, host, (size_t)n * 4, cudaMemcpyHostToDevice));
  return 0;
}
int cib_engine_download(cib_engine_t e_, int which, float* host, long long n) {
  Engine* e = ENG(e_); float* d = which_buf(e, which);
  if (!d || n > e->nparam) return fail("download: bad buffer or size");
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  CIB_CUDA(cudaMemcpy(host, d, (size_t)n * 4, cudaMemcpyDeviceToHost));
  return 0;
}
int cib_engine_set_hyper(cib_engine_t e_, const CibHyper* h) {
  Engine* e = ENG(e_);
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  HyperBlock hb;
  CIB_CUDA(cudaMemcpy(&hb, e->hb, sizeof hb, cudaMemcpyDeviceToHost));
  hb.lr = h->lr; hb.beta1 = h->beta1; hb.beta2 = h->beta2; hb.eps = h->eps; hb.rho = h->rho; hb.momentum = h->momentum;
  hb.schedule_decay = h->schedule_decay; hb.decay = h->decay;
  hb.m_schedule[0] = hb.m_schedule[1] = h->m_schedule;
  hb.iterations = h->iterations; hb.seed = h->seed; hb.training = 1; hb.ticket = 0;
  CIB_CUDA(cudaMemcpy(e->hb, &hb, sizeof hb, cudaMemcpyHostToDevice));
  return 0;
}
int cib_engine_set_lr(cib_engine_t e_, float lr) {
  Engine* e = ENG(e_);
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  CIB_CUDA(cudaMemcpy(&e->hb->lr, &lr, 4, cudaMemcpyHostToDevice));
  return 0;
}
int cib_engine_get_metrics(cib_engine_t e_, CibMetrics* out) {
  Engine* e = ENG(e_);
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  HyperBlock hb;
  CIB_CUDA(cudaMemcpy(&hb, e->hb, sizeof hb, cudaMemcpyDeviceToHost));
  out->loss_sum = hb.loss_sum; out->correct_sum = hb.correct_sum; out->sample_sum = hb.sample_sum;
  out->step_loss = hb.step_loss; out->step_acc = hb.step_acc; out->iterations = hb.iterations;
  out->m_schedule = hb.m_schedule[hb.iterations & 1u]; out->batch_cursor = hb.batch_cursor;
  return check_peer_error(e);
}
int cib_engine_reset_metrics(cib_engine_t e_) {
  Engine* e = ENG(e_);
  CIB_CUDA(cudaMemsetAsync(&e->hb->loss_sum, 0, 3 * sizeof(float), e->stream));
  return 0;
}
static int push_dataset(Engine* e) {
  struct { unsigned int cursor; int n_total; const float* x; const float* y; const int* perm; } d;
  d.cursor = 0; d.n_total = e->n_data; d.x = e->data_x; d.y = e->data_y; d.perm = e->has_perm ? e->perm : nullptr;
  static_assert(sizeof(d) == 32, "layout must match the tail of HyperBlock");
  static_assert(offsetof(HyperBlock, batch_cursor) % 8 == 0 && offsetof(HyperBlock, data_x) == offsetof(HyperBlock, batch_cursor) + 8 &&
                offsetof(HyperBlock, perm) + 8 == sizeof(HyperBlock), "HyperBlock dataset block layout");
  CIB_CUDA(cudaMemcpyAsync(&e->hb->batch_cursor, &d, sizeof d, cudaMemcpyHostToDevice, e->stream));
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  return 0;
}
int cib_engine_set_dataset(cib_engine_t e_, const float* dev_x, const float* dev_y, int n, int label_w) {
  Engine* e = ENG(e_);
  if (label_w != e->label_w) return fail("set_dataset: label width does not match the model head");
  e->data_x = dev_x; e->data_y = dev_y; e->n_data = n; e->has_perm = false;
  return push_dataset(e);
}
int cib_engine_set_perm(cib_engine_t e_, const int* host_perm, int n) {
  Engine* e = ENG(e_);
  if (n != e->n_data) return fail("set_perm: permutation length != dataset length");
  if (host_perm) {
    CIB_CUDA(cudaStreamSynchronize(e->stream));
    if (e->perm_cap < n) {
      cudaFree(e->perm); e->perm = nullptr;
      CIB_CUDA(cudaMalloc(&e->perm, (size_t)n * 4)); e->perm_cap = n;
    }
    CIB_CUDA(cudaMemcpy(e->perm, host_perm, (size_t)n * 4, cudaMemcpyHostToDevice));
    e->has_perm = true;
  } else {
    e->has_perm = false;
  }
  return push_dataset(e);
}
int cib_engine_train_steps(cib_engine_t e_, int nsteps) {
  Engine* e = ENG(e_);
  if (!e->data_x) return fail("train_steps: no dataset set");
  NvtxRange r("cib.train_steps");
  CIB_TRY(e->build_graphs());
  for (int i = 0; i < nsteps; ++i) {
    if (e->use_graph) CIB_CUDA(cudaGraphLaunch(e->g_train, e->stream));
    else { e->kcount = 0; CIB_TRY(e->enqueue_step(e->stream, 0, true, true)); e->train_kernels = e->kcount; }
  }
  return 0;
}
// Device-timed replay: CUDA events on the engine stream around exactly `nsteps` train steps (synchronised on both sides).
int cib_engine_time_steps(cib_engine_t e_, int nsteps, float* ms_out) {
  Engine* e = ENG(e_);
  if (!e->data_x) return fail("time_steps: no dataset set");
  CIB_TRY(e->build_graphs());
  cudaEvent_t a, b;
  CIB_CUDA(cudaEventCreate(&a)); CIB_CUDA(cudaEventCreate(&b));
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  CIB_CUDA(cudaEventRecord(a, e->stream));
  for (int i = 0; i < nsteps; ++i) {
    if (e->use_graph) CIB_CUDA(cudaGraphLaunch(e->g_train, e->stream));
    else CIB_TRY(e->enqueue_step(e->stream, 0, true, true));
  }
  CIB_CUDA(cudaEventRecord(b, e->stream));
  CIB_CUDA(cudaEventSynchronize(b));
  CIB_CUDA(cudaEventElapsedTime(ms_out, a, b));
  cudaEventDestroy(a); cudaEventDestroy(b);
  return 0;
}
// The same with `preroll` untimed steps issued right before the first event, with no host synchronisation in between: under
// data parallelism the preroll step's collective lines the ranks up ON THE DEVICE, so the timed region does not contain the
// host-side start skew between processes (which a 20-step region would otherwise mostly consist of).
int cib_engine_time_steps_aligned(cib_engine_t e_, int preroll, int nsteps, float* ms_out) {
  Engine* e = ENG(e_);
  if (!e->data_x) return fail("time_steps: no dataset set");
  CIB_TRY(e->build_graphs());
  cudaEvent_t a, b;
  CIB_CUDA(cudaEventCreate(&a)); CIB_CUDA(cudaEventCreate(&b));
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  for (int i = 0; i < preroll + nsteps; ++i) {
    if (i == preroll) CIB_CUDA(cudaEventRecord(a, e->stream));
    if (e->use_graph) CIB_CUDA(cudaGraphLaunch(e->g_train, e->stream));
    else CIB_TRY(e->enqueue_step(e->stream, 0, true, true));
  }
  CIB_CUDA(cudaEventRecord(b, e->stream));
  CIB_CUDA(cudaEventSynchronize(b));
  CIB_CUDA(cudaEventElapsedTime(ms_out, a, b));
  cudaEventDestroy(a); cudaEventDestroy(b);
  return 0;
}
// 1: the batch cursor wraps to 0 instead of running past the end of the bound dataset (steady-state timing loops that replay
// more steps than one pass over the data); 0 (default): an epoch ends where the data ends (Keras fit semantics).
int cib_engine_set_wrap(cib_engine_t e_, int on) {
  Engine* e = ENG(e_);
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  unsigned int v = on ? 1u : 0u;
  CIB_CUDA(cudaMemcpy(&e->hb->cursor_wrap, &v, 4, cudaMemcpyHostToDevice));
  return 0;
}
int cib_engine_eval_steps(cib_engine_t e_, int nsteps, float* dev_probs_out) {
  Engine* e = ENG(e_);
  if (!e->data_x) return fail("eval_steps: no dataset set");
  NvtxRange r("cib.eval_steps");
  CIB_TRY(e->build_graphs());
  for (int i = 0; i < nsteps; ++i) {
    if (e->use_graph) CIB_CUDA(cudaGraphLaunch(e->g_eval, e->stream));
    else { e->kcount = 0; CIB_TRY(e->enqueue_step(e->stream, 0, false, true)); e->eval_kernels = e->kcount; }
    if (dev_probs_out)
      CIB_CUDA(cudaMemcpyAsync(dev_probs_out + (size_t)i * e->B * e->label_w, e->probs, (size_t)e->B * e->label_w * 4,
                               cudaMemcpyDeviceToDevice, e->stream));
  }
  return 0;
}
int cib_engine_train_batch_host(cib_engine_t e_, const float* host_x, const float* host_y, int n_valid, float* host_loss_out) {
  Engine* e = ENG(e_);
  NvtxRange r("cib.train_batch_host");
  CIB_TRY(e->build_graphs());
  const int k = (int)(e->host_calls & 1);
  const size_t xbytes = (size_t)e->B * e->in_h * e->in_w * e->in_c * 4, ybytes = (size_t)e->B * e->label_w * 4;
  // buffer k was last read by the step issued two calls ago
  if (e->host_calls >= 2) CIB_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_done[k], 0));
  // the input goes up in four pieces on four streams (see Engine::copy_more); small inputs in one piece
  const int nch = xbytes >= (1u << 20) ? g_copy_chunks : 1;
  const size_t chunk = ((xbytes / nch) + 255) & ~(size_t)255;
  for (int c = 1; c < nch; ++c) {
    const size_t off = (size_t)c * chunk;
    if (off >= xbytes) break;
    cudaStream_t cs = e->copy_more[c - 1];
    if (e->host_calls >= 2) CIB_CUDA(cudaStreamWaitEvent(cs, e->ev_done[k], 0));
    CIB_CUDA(cudaMemcpyAsync((char*)e->xin[k] + off, (const char*)host_x + off, std::min(chunk, xbytes - off), cudaMemcpyHostToDevice, cs));
    CIB_CUDA(cudaEventRecord(e->ev_copy_more[k][c - 1], cs));
    CIB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_copy_more[k][c - 1], 0));
  }
  CIB_CUDA(cudaMemcpyAsync(e->xin[k], host_x, std::min(chunk, xbytes), cudaMemcpyHostToDevice, e->copy_stream));
  CIB_CUDA(cudaMemcpyAsync(e->yb[k], host_y, ybytes, cudaMemcpyHostToDevice, e->copy_stream));
  CIB_CUDA(cudaEventRecord(e->ev_copy[k], e->copy_stream));
  CIB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_copy[k], 0));
  if (n_valid != e->B || e->host_calls == 0) {
    set_sample_weights_kernel<<<(e->B + 127) / 128, 128, 0, e->stream>>>(e->sw, e->B, n_valid);
    CIB_CUDA(cudaGetLastError());
  }
  if (e->use_graph) CIB_CUDA(cudaGraphLaunch(e->g_train_host[k], e->stream));
  else { e->kcount = 0; CIB_TRY(e->enqueue_step(e->stream, k, true, false)); e->train_kernels = e->kcount; }
  CIB_CUDA(cudaEventRecord(e->ev_done[k], e->stream));
  // host_loss_out != NULL: classic copy node after the step.  NULL: the loss head has already written {loss, acc} into the engine's
  // pinned host ring (cib_engine_loss_ring) -- nothing sits between two step graphs on the stream.
  if (host_loss_out)
    CIB_CUDA(cudaMemcpyAsync(host_loss_out, &e->hb->step_loss, 2 * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
  e->host_calls++;
  return 0;
}
// Pinned host ring {loss, accuracy} x entries written by the loss head of every host-fed training step; slot = (optimizer step
// counter before the step) & (entries - 1).  Valid after cib_engine_sync; read at least once every `entries` steps.
int cib_engine_loss_ring(cib_engine_t e_, float** host_ring, int* entries) {
  Engine* e = ENG(e_);
  if (host_ring) *host_ring = e->loss_pin;
  if (entries) *entries = CIB_LOSS_RING;
  return 0;
}
int cib_engine_sync(cib_engine_t e_) {
  Engine* e = ENG(e_);
  CIB_CUDA(cudaStreamSynchronize(e->copy_stream));
  for (int c = 0; c < 7; ++c) CIB_CUDA(cudaStreamSynchronize(e->copy_more[c]));
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  return check_peer_error(e);
}
// CIB_DEBUG_BOUNDS=1 only: *bad = number of canary bytes around the engine's buffers that were overwritten (0 = intact); the name of
// the first damaged buffer goes to cib_last_error().  Synchronises the engine.
int cib_engine_check_canaries(cib_engine_t e_, long long* bad) {
  Engine* e = ENG(e_);
  CIB_CUDA(cudaStreamSynchronize(e->stream)); CIB_CUDA(cudaStreamSynchronize(e->side)); CIB_CUDA(cudaStreamSynchronize(e->upd));
  if (!g_debug_bounds) { *bad = -1; return 0; }
  std::string where;
  CIB_TRY(e->check_canaries(bad, &where));
  if (*bad) g_last_error = "canary damaged next to: " + where;
  return 0;
}
int cib_engine_launch_count(cib_engine_t e_, int* t, int* v) {
  Engine* e = ENG(e_);
  if (t) *t = e->train_kernels; if (v) *v = e->eval_kernels;
  return 0;
}
int cib_engine_buffer(cib_engine_t e_, int which, void** ptr, long long* nfloats) {
  Engine* e = ENG(e_);
  void* p = nullptr; long long n = 0;
  if (which >= 0 && which <= 3) { p = which_buf(e, which); n = e->nparam; }
  else if (which == 4) { p = e->xin[0]; n = (long long)e->B * e->in_h * e->in_w * e->in_c; }
  else if (which == 5) { p = e->yb[0]; n = (long long)e->B * e->label_w; }
  else if (which == 6) { p = e->probs; n = (long long)e->B * e->label_w; }
  else if (which == 7) { p = e->sw; n = e->B; }
  else if (which >= 100 && which < 100 + (int)e->blocks.size()) { p = e->blocks[which - 100].act; n = e->blocks[which - 100].act_n; }
  else if (which >= 200 && which < 200 + (int)e->blocks.size()) { p = e->blocks[which - 200].dz; n = e->blocks[which - 200].dz_n; }
  else return fail("buffer: unknown id");
  *ptr = p; if (nfloats) *nfloats = n;
  return 0;
}
int cib_engine_forward_backward(cib_engine_t e_, int training, int backward) {
  Engine* e = ENG(e_);
  e->kcount = 0;
  CIB_TRY(e->enqueue_forward(e->stream, 0, training != 0, false, nullptr));
  if (training && backward) CIB_TRY(e->enqueue_backward(e->stream, 0));
  CIB_CUDA(cudaStreamSynchronize(e->side));
  CIB_CUDA(cudaStreamSynchronize(e->stream));
  return 0;
}
// Several trials share one GPU (engine --slots k): a trial marked high gets its compute streams re-created at the greatest stream
// priority, so its kernels are picked first whenever the SMs have room (stream priorities between packed trials).  The
// graphs are captured again at the next step because captured kernel nodes inherit the priority of the stream they were captured from.
int cib_engine_set_stream_priority(cib_engine_t e_, int high) {
  Engine* e = ENG(e_);
  std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
  int least = 0, greatest = 0;
  CIB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
  const int prio = high ? greatest : least;
  cudaStream_t* all[2] = {&e->stream, &e->side};            // `upd` (the early bucket) always has the greatest priority
  for (cudaStream_t* s : all) {
    if (*s) { CIB_CUDA(cudaStreamSynchronize(*s)); CIB_CUDA(cudaStreamDestroy(*s)); *s = nullptr; }
    CIB_CUDA(cudaStreamCreateWithPriority(s, cudaStreamNonBlocking, prio));
  }
  e->stream_priority = prio;
  e->graphs_dirty = true;
  return 0;
}
int cib_engine_get_stream_priority(cib_engine_t e_, int* prio, int* least_out, int* greatest_out) {
  Engine* e = ENG(e_);
  int least = 0, greatest = 0;
  CIB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
  int p = 0;
  CIB_CUDA(cudaStreamGetPriority(e->stream, &p));
  if (prio) *prio = p; if (least_out) *least_out = least; if (greatest_out) *greatest_out = greatest;
  return 0;
}

}  // extern "C"

#include "comm_api.cuh"
#include "nvls_api.cuh"
#include "ops_api.cuh"

// =============================================================================================================
// Allocation-free variant of the fused dense-layer forward.
//
// cib_dense_relu_fwd() needs a caller-provided fp32 scratch matrix for its split-K partial sums.  Some callers of this
// library have no way to provide one at the moment they call:
//   * HPO engines pack several trials on one GPU, each replaying CUDA graphs on its own stream.  cudaMalloc / cudaFree
//     synchronise the whole device, so one trial asking for a temporary buffer would stall every other trial; such a
//     trial must not allocate once it is running.
//   * The cluster scheduler hands every trial a fixed slice of HBM up front.  A helper that allocates "just a little"
//     behind its back can push a neighbour over its slice and fail in the middle of a run.
// For them the scratch lives in the file-scope array below: 256 KiB that every GPU which loads this library holds until
// the process exits.  That fixes the largest problem the variant accepts: M * N <= CIB_NOALLOC_MAX_FLOATS.
//
// Things a caller has to know:
//   * The first call on a device does a one-time set-up for that device (kernel attributes and the device address of
//     the scratch array).  Make that first call outside of any stream capture.
//   * Every call uses the same array.  Calls issued to the same stream run one after the other and are fine; calls that
//     can overlap on the GPU (different streams, different host threads) race on the scratch and give wrong results.
//   * The array keeps whatever the previous call left in it, so each call clears the part it is about to use first.
// =============================================================================================================
#define CIB_NOALLOC_MAX_FLOATS (256 * 256)
#define CIB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) float g_noalloc_scratch[CIB_NOALLOC_MAX_FLOATS];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device address: the address a kernel may use is
// obtained per device with cudaGetSymbolAddress and cached here.
static float* g_noalloc_dev_ptr[CIB_NOALLOC_MAX_DEVICES] = {nullptr};

extern "C" int cib_dense_relu_fwd_noalloc(const float* A, const float* W, const float* bias, float* out, int M, int N, int K,
                                          void* stream) {
  if (M < 1 || N < 1 || K < 1) return fail("dense_relu_fwd_noalloc: M, N, K must be positive");
  if ((long long)M * N > (long long)CIB_NOALLOC_MAX_FLOATS)
    return fail("dense_relu_fwd_noalloc: M * N = " + std::to_string((long long)M * N) + " floats exceeds the built-in scratch of " +
                std::to_string(CIB_NOALLOC_MAX_FLOATS) + " floats (256 KiB); use cib_dense_relu_fwd with your own scratch buffer");
  int dev = -1;
  CIB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= CIB_NOALLOC_MAX_DEVICES) return fail("dense_relu_fwd_noalloc: unsupported device ordinal");
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    // one-time set-up for this device, with this device current
    cudaDeviceProp prop;
    CIB_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) return fail("cib200 needs an sm_100 GPU");
    g_num_sms = prop.multiProcessorCount;
    if (g_device < 0) g_device = dev;
    CIB_TRY(ensure_driver());
    CIB_TRY(init_kernel_attributes());
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(std::string("dense_relu_fwd_noalloc: cudaGetSymbolAddress(scratch) -> ") + cudaGetErrorString(e));
    g_noalloc_dev_ptr[dev] = static_cast<float*>(p);
  }
  float* z = g_noalloc_dev_ptr[dev];
  cudaStream_t st = (cudaStream_t)stream;
  // the split-K partial sums are accumulated with red.add: the slice used by this call has to start at zero
  CIB_CUDA(cudaMemsetAsync(z, 0, (size_t)M * N * sizeof(float), st));
  return dense_relu_fwd_impl(A, W, bias, out, z, M, N, K, st);
}
