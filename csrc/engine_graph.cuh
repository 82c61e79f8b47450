// Optimizer / fused all-reduce enqueue, whole-step sequencing and CUDA-graph capture.
#pragma once
#include "engine_steps.cuh"

namespace cib {

static int ar_grid_for(long long n4) {
  // enough CTAs to pull peer memory at full NVLink bandwidth, never more than the flag table holds
  long long g = (n4 + 2047) / 2048;
  if (g < 1) g = 1;
  return (int)std::min<long long>(g, std::min(CIB_AR_MAX_CTAS, g_num_sms));
}

static int fill_commdev(CommDev* cd, const Comm* c, int rank, size_t off_grads, size_t off_params, size_t off_flags) {
  memset(cd, 0, sizeof *cd);
  cd->rank = rank; cd->world = c->world;
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

int Engine::enqueue_update_range(cudaStream_t st, long long start, long long count, bool advance) {
  if (count <= 0) return 0;
  const int n4 = (int)(count / 4);
  if (comm && comm->world > 1) {
    AllreduceOptParams p{};
    CIB_TRY(fill_commdev(&p.comm, comm, comm_rank, off_grads + (size_t)start * 4, off_params + (size_t)start * 4, off_flags));
    p.slot0 = slot0 + start; p.slot1 = opt_kind == 0 ? nullptr : slot1 + start; p.n4 = n4; p.kind = opt_kind;
    const long long bytes = count * 4;
    p.one_shot = ar_mode == 1 ? 1 : ar_mode == 2 ? 0 : (bytes <= 256 * 1024 ? 1 : 0);
    p.hb = hb; p.epoch_ctr = epoch_ctr; p.advance = advance ? 1 : 0;
    allreduce_opt_kernel<<<ar_grid_for(n4), 512, 0, st>>>(p); ++kcount;
  } else {
    optimizer_kernel<<<grid_for(n4, 256, 4), 256, 0, st>>>(params + start, grads + start, slot0 + start,
                                                           opt_kind == 0 ? nullptr : slot1 + start, n4, opt_kind, hb, advance ? 1 : 0); ++kcount;
  }
  CIB_CUDA(cudaGetLastError());
  return 0;
}

int Engine::enqueue_update(cudaStream_t st) { return enqueue_update_range(st, 0, nparam, true); }

int Engine::enqueue_step(cudaStream_t st, int k, bool training, bool gather) {
  CIB_TRY(enqueue_forward(st, k, training, gather, nullptr));
  if (training) {
    const long long split = g_early_update ? split_at() : 0;
    CIB_TRY(enqueue_backward(st, k, split > 0));
    CIB_TRY(enqueue_update_range(st, 0, split > 0 ? split : nparam, true));
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
  e = cudaGraphInstantiate(out, g, 0);
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
    for (int k = 0; k < 2; ++k) { kcount = 0; CIB_TRY(capture(&g_train_host[k], k, true, false)); train_kernels = kcount; }
  }
  graphs_dirty = false;
  return 0;
}

}  // namespace cib
