// Optimizer / fused all-reduce enqueue, whole-step sequencing and CUDA-graph capture.
#pragma once
#include "engine_steps.cuh"

namespace cib {

static int g_par_ctas = 32;            // CTAs of the push all-reduce (<= CIB_PAR_MAX_CTAS); CIB_PAR_CTAS overrides

static int ar_grid_for(long long n4) {
  // enough CTAs to pull peer memory at full NVLink bandwidth, never more than the flag table holds
  long long g = (n4 + 2047) / 2048;
  if (g < 1) g = 1;
  return (int)std::min<long long>(g, std::min(CIB_AR_MAX_CTAS, g_num_sms));
}

static int fill_commdev(CommDev* cd, const Comm* c, int rank, size_t off_grads, size_t off_params, size_t off_flags) {
  memset(cd, 0, sizeof *cd);
  cd->rank = rank; cd->world = c->world; cd->timeout_ns = g_peer_timeout_ns;
  for (int q = 0; q < c->world; ++q) {
    uint8_t* base = c->bases[c->ipc ? 0 : rank][q];
    cd->grads[q] = (float*)(base + off_grads);
    cd->params[q] = (float*)(base + off_params);
    cd->flags[q] = (unsigned int*)(base + off_flags);
  }
  return 0;
}

// The flat parameter / gradient buffers are updated in two buckets ("dense-layer bucket launched while conv
// backward still runs"): bucket A = the dense stack (96 % of the RPV model's parameters; its gradients are the FIRST to be
// complete in backward) is all-reduced + updated on the side stream while the convolution backward runs on the main stream,
// bucket B = the convolution stack at the end of the step.  Only bucket B moves the optimizer's step counter on.
long long Engine::split_at() const {
  for (size_t i = 0; i < blocks.size(); ++i)
    if (blocks[i].b.kind != 0) return i == 0 ? 0 : (long long)blocks[i].b.w_off;
  return 0;
}

// one-shot (every rank reduces and updates the whole bucket) or two-shot (each rank owns a slice); ar_mode 1 / 2 force it
bool Engine::bucket_one_shot(long long count) const {
  bool one = ar_mode == 1 ? true : ar_mode == 2 ? false : count * 4 <= 256 * 1024;
  if ((size_t)(count / 4) > oneshot_cap4) one = false;
  return one;
}

// small one-shot buckets use the LL protocol (data carry their own flags): CIB_AR_LL=0 / cib_set_ar_ll(0) switches it off
static int g_ar_ll = 1;
bool Engine::bucket_ll(long long count) const {
  return g_ar_ll && bucket_one_shot(count) && count / 4 <= CIB_LL_MAX4;
}

// After two-shot updates every rank holds only its own slice of the optimizer moments: pull the other slices from their owners.
// Collective: every rank calls it at the same point.  (The pull kernel's two-shot mode shards per CTA, not per rank -- it is an
// A/B switch for measurements only, not to be combined with checkpoints.)
int Engine::gather_slots() {
  if (!comm || comm->world < 2) return 0;
  const long long split = use_early_update() ? split_at() : 0;
  const long long starts[2] = {0, split}, counts[2] = {split > 0 ? split : nparam, split > 0 ? nparam - split : 0};
  for (int b = 0; b < 2; ++b) {
    if (counts[b] <= 0 || bucket_one_shot(counts[b])) continue;
    GatherSlotsParams p{};
    CIB_TRY(fill_commdev(&p.comm, comm, comm_rank, off_slot0 + (size_t)starts[b] * 4, off_slot1 + (size_t)starts[b] * 4, off_flags));
    if (opt_kind == 0) for (int q = 0; q < CIB_MAX_RANKS; ++q) p.comm.params[q] = nullptr;
    p.n4 = (int)(counts[b] / 4); p.S4 = (p.n4 + comm->world - 1) / comm->world;
    p.epoch_ctr = epoch_ctr; p.ticket = epoch_ctr + 1;
    gather_slots_kernel<<<ar_grid_for(p.n4), 512, 0, stream>>>(p);
    CIB_CUDA(cudaGetLastError());
  }
  return 0;
}

// 0 (default) = push kernel, 1 = the older pull kernel (two barriers, peer loads); CIB_AR_PULL=1 or cib_set_ar_pull(1), for A/B runs
static int g_ar_pull = 0;

int Engine::enqueue_update_range(cudaStream_t st, long long start, long long count, bool advance) {
  if (count <= 0) return 0;
  const int n4 = (int)(count / 4);
  if (comm && comm->world > 1 && g_ar_pull) {
    AllreduceOptParams p{};
    CIB_TRY(fill_commdev(&p.comm, comm, comm_rank, off_grads + (size_t)start * 4, off_params + (size_t)start * 4, off_flags));
    p.slot0 = slot0 + start; p.slot1 = opt_kind == 0 ? nullptr : slot1 + start; p.n4 = n4; p.kind = opt_kind;
    const long long bytes = count * 4;
    p.one_shot = ar_mode == 1 ? 1 : ar_mode == 2 ? 0 : (bytes <= 256 * 1024 ? 1 : 0);
    p.hb = hb; p.epoch_ctr = epoch_ctr; p.advance = advance ? 1 : 0;
    allreduce_opt_kernel<<<ar_grid_for(n4), 512, 0, st>>>(p); ++kcount;
  } else if (comm && comm->world > 1 && nv && nv->ready && !bucket_one_shot(count)) {
    // large bucket, gradients bound to the multicast object: in-switch reduction
    const int W = comm->world;
    NvlsArParams p{};
    CIB_TRY(fill_commdev(&p.comm, comm, comm_rank, off_grads, off_params + (size_t)start * 4, off_flags));
    p.mc_grads = (const float*)nv->mcva + start; p.grads_l = grads + start;
    p.params_l = params + start; p.slot0 = slot0 + start; p.slot1 = opt_kind == 0 ? nullptr : slot1 + start;
    p.n4 = n4; p.S4 = (n4 + W - 1) / W; p.kind = opt_kind; p.advance = advance ? 1 : 0;
    p.hb = hb; p.epoch_ctr = epoch_ctr; p.ticket = (unsigned int*)(arena + off_ctr2) + 8;
    const long long gmax = advance ? CIB_PAR_MAX_CTAS : g_par_ctas;
    const int G = (int)std::max<long long>(1, std::min<long long>(gmax, ((long long)p.S4 + CIB_PAR_THREADS - 1) / CIB_PAR_THREADS));
    cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(G); cfg.blockDim = dim3(CIB_PAR_THREADS); cfg.stream = st;
    cudaLaunchAttribute at[1];
    int least = 0, greatest = 0;
    CIB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    at[0].id = cudaLaunchAttributePriority; at[0].val.priority = greatest;
    cfg.attrs = at; cfg.numAttrs = 1;
    CIB_CUDA(cudaLaunchKernelEx(&cfg, nvls_allreduce_opt_kernel, p));
    ++kcount;
  } else if (comm && comm->world > 1) {
    const int W = comm->world, R = comm_rank;
    const int bucket = start > 0 ? 1 : 0;
    PushArParams p{};
    p.one_shot = bucket_one_shot(count) ? 1 : 0;
    // receive areas of the two buckets: bucket 0 first.  In float4 units a two-shot bucket takes W x S4, a one-shot bucket
    // 2 x W x n4 (flag protocol) or 4 x W x n4 (LL protocol: 32 bytes per float4, two parities)
    const long long split4 = bucket == 1 ? start / 4 : 0;
    size_t roff4 = 0;
    if (bucket == 1) roff4 = bucket_one_shot(start) ? (size_t)(bucket_ll(start) ? 4 : 2) * W * split4 : (size_t)W * ((split4 + W - 1) / W);
    const bool ll = p.one_shot && bucket_ll(count);
    p.S4 = p.one_shot ? n4 : (n4 + W - 1) / W;
    const size_t need4 = roff4 + (p.one_shot ? (size_t)(ll ? 4 : 2) * W * n4 : (size_t)W * p.S4);
    if (need4 > recv_cap4) return fail("engine: receive area too small for this all-reduce configuration");
    p.grads = grads + start; p.params_l = params + start; p.slot0 = slot0 + start; p.slot1 = opt_kind == 0 ? nullptr : slot1 + start;
    for (int q = 0; q < W; ++q) {
      uint8_t* base = comm->bases[comm->ipc ? 0 : R][q];
      p.recv[q] = (float*)(base + off_recv) + 4 * roff4;
      p.params[q] = (float*)(base + off_params) + start;
      p.flags[q] = (unsigned int*)(base + off_flags2) + (size_t)bucket * 2 * CIB_PAR_MAX_CTAS * CIB_MAX_RANKS;
    }
    p.rank = R; p.world = W; p.n4 = n4; p.kind = opt_kind; p.advance = advance ? 1 : 0;
    p.hb = hb;
    p.epoch_ctr = (unsigned int*)(arena + off_ctr2) + bucket * 2; p.ticket = p.epoch_ctr + 1;
    p.timeout_ns = g_peer_timeout_ns;
    p.trace = ar_trace ? ar_trace + bucket * 8 : nullptr;
    // the early bucket shares the SMs with the convolution backward: few CTAs; the step's last bucket has the GPU to itself
    const long long gmax = advance ? CIB_PAR_MAX_CTAS : g_par_ctas;
    const int G = (int)std::max<long long>(1, std::min<long long>(gmax, ((long long)p.S4 + CIB_PAR_THREADS - 1) / CIB_PAR_THREADS));
    // Explicit launch priority (recorded in the captured kernel node): the block scheduler serves grids of equal priority in order, so
    // without it the early bucket's CTAs queue behind the ~150 not-yet-placeable CTAs of the convolution wgrad launched just before
    // (measured: 25 us late); with the greatest priority they slip onto the SMs next to the running convolution CTAs at once.
    cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(G); cfg.blockDim = dim3(CIB_PAR_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    int least = 0, greatest = 0;
    CIB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    at[0].id = cudaLaunchAttributePriority; at[0].val.priority = greatest;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (ll) CIB_CUDA(cudaLaunchKernelEx(&cfg, ll_allreduce_opt_kernel, p));
    else CIB_CUDA(cudaLaunchKernelEx(&cfg, push_allreduce_opt_kernel, p));
    ++kcount;
  } else {
    optimizer_kernel<<<grid_for(n4, 256, g_opt_grid_mult), 256, 0, st>>>(params + start, grads + start, slot0 + start,
                                                           opt_kind == 0 ? nullptr : slot1 + start, n4, opt_kind, hb, advance ? 1 : 0); ++kcount;
  }
  CIB_CUDA(cudaGetLastError());
  return 0;
}

int Engine::enqueue_update(cudaStream_t st) { return enqueue_update_range(st, 0, nparam, true); }

// two buckets per step?  g_early_update: 0 never, 1 under data parallelism (the early bucket hides the dense all-reduce behind the
// convolution backward; on one GPU the extra launch costs more than it hides, measured 0.228 vs 0.224 ms/step), 2 always
bool Engine::use_early_update() const {
  return g_early_update == 2 || (g_early_update == 1 && comm && comm->world > 1);
}

int Engine::enqueue_step(cudaStream_t st, int k, bool training, bool gather) {
  if (ar_trace && training) { stamp_kernel<<<1, 1, 0, st>>>(ar_trace + 16); }
  CIB_TRY(enqueue_forward(st, k, training, gather, nullptr));
  if (training) {
    const long long split = use_early_update() ? split_at() : 0;
    CIB_TRY(enqueue_backward(st, k, split > 0));
    if (ar_trace) { stamp_kernel<<<1, 1, 0, st>>>(ar_trace + 19); }
    CIB_TRY(enqueue_update_range(st, 0, split > 0 ? split : nparam, true));
    if (ar_trace) { stamp_kernel<<<1, 1, 0, st>>>(ar_trace + 20); }
  }
  return 0;
}

int Engine::capture(cudaGraphExec_t* out, int k, bool training, bool gather) {
  if (*out) { cudaGraphExecDestroy(*out); *out = nullptr; }
  cudaGraph_t g = nullptr;
  CIB_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
  int rc = enqueue_step(stream, k, training, gather);
  cudaError_t e = cudaStreamEndCapture(stream, &g);
  if (rc) { if (g) cudaGraphDestroy(g); return rc; }
  if (e != cudaSuccess) return fail(std::string("cudaStreamEndCapture -> ") + cudaGetErrorString(e));
  // per-node priorities: without this flag every node runs at the priority of the stream the graph is launched into, and the early
  // bucket's high-priority launch attribute would be ignored
  e = cudaGraphInstantiate(out, g, cudaGraphInstantiateFlagUseNodePriority);
  cudaGraphDestroy(g);
  if (e != cudaSuccess) return fail(std::string("cudaGraphInstantiate -> ") + cudaGetErrorString(e));
  return 0;
}

int Engine::build_graphs() {
  if (!graphs_dirty) return 0;
  NvtxRange r("cib.build_graphs");
  std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
  CIB_CUDA(cudaStreamSynchronize(stream));
  // warm every kernel once outside capture so cudaFuncSetAttribute / lazy module loading never lands in a capture
  kcount = 0;
  if (use_graph) {
    kcount = 0; CIB_TRY(capture(&g_train, 0, true, true)); train_kernels = kcount;
    kcount = 0; CIB_TRY(capture(&g_eval, 0, false, true)); eval_kernels = kcount;
    // the host-fed graphs have no gather kernel: counted separately so that launch_count reports the device-resident step
    for (int k = 0; k < 2; ++k) { kcount = 0; CIB_TRY(capture(&g_train_host[k], k, true, false)); host_kernels = kcount; }
  }
  graphs_dirty = false;
  return 0;
}

}  // namespace cib
