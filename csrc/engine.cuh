// The train-step executor: plans every buffer once, records forward + backward + (all-reduce +) optimizer into
// CUDA graphs, and replays them per batch with no host synchronisation inside an epoch.
// Replaces Keras' fit_loop / train_function / test_function (reference rpv.py:103-106, DistHPO_mnist.ipynb:187-190).
#pragma once
#include "runtime.cuh"
#include "../include/cib200.h"

namespace cib {

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Comm;

struct BlockRt {                 // runtime state of one fused block
  CibBlock b;
  float* act = nullptr;          // stored activation of this block (conv: [B][Hs][Ws][Co]; dense: [B][N])
  float* dz = nullptr;           // gradient w.r.t. this block's pre-activation (conv: pre-pool grid, or the
                                 // activation grid when `from_pooled`; dense: [B][N])
  float* z = nullptr;            // dense: split-K accumulator [B][N]
  long long act_n = 0, dz_n = 0;
  int Hs = 0, Ws = 0;            // conv: stored grid
  bool from_pooled = false;      // conv block 0 fast path: wgrad consumes the masked gradient at activation grid
};

// CIB_DEBUG_BOUNDS=1 (compute-sanitizer is not available on this GPU pool): every engine buffer is allocated with a 256-byte canary
// band in front of and behind it, filled with 0xCB; cib_engine_check_canaries verifies all bands.  A kernel that writes outside
// its buffer -- the tensor-map / halo-loader class of bug -- is caught at the next check instead of silently corrupting a neighbour.
static int g_debug_bounds = 0;
static constexpr size_t CIB_CANARY = 256;
static constexpr int CIB_LOSS_RING = 4096;     // entries of the pinned host ring the loss head writes {loss, acc} of host-fed steps into
struct GuardedAlloc { uint8_t* raw; size_t bytes; const char* what; };

struct Engine {
  std::vector<GuardedAlloc> guarded;
  template <typename T>
  int dev_alloc(T** out, size_t bytes, const char* what) {
    if (!g_debug_bounds) { CIB_CUDA(cudaMalloc(out, bytes)); return 0; }
    const size_t padded = (bytes + 255) & ~(size_t)255;
    uint8_t* raw = nullptr;
    CIB_CUDA(cudaMalloc(&raw, padded + 2 * CIB_CANARY));
    CIB_CUDA(cudaMemset(raw, 0xCB, padded + 2 * CIB_CANARY));
    guarded.push_back({raw, padded, what});
    *out = reinterpret_cast<T*>(raw + CIB_CANARY);
    return 0;
  }
  void dev_free(void* p) {
    if (!p) return;
    if (g_debug_bounds) {
      for (auto& g : guarded) if (g.raw + CIB_CANARY == p) { cudaFree(g.raw); g.raw = nullptr; return; }
    }
    cudaFree(p);
  }
  // number of canary bytes that no longer hold the pattern; `where` names the first damaged buffer
  int check_canaries(long long* bad_out, std::string* where) {
    long long bad = 0;
    std::vector<uint8_t> host(CIB_CANARY);
    for (auto& g : guarded) {
      if (!g.raw) continue;
      // the band behind a buffer starts at its UNPADDED end rounded up to 256 bytes: writes into the padding itself are not seen
      for (int side = 0; side < 2; ++side) {
        CIB_CUDA(cudaMemcpy(host.data(), side ? g.raw + CIB_CANARY + g.bytes : g.raw, CIB_CANARY, cudaMemcpyDeviceToHost));
        long long b = 0;
        for (uint8_t v : host) b += v != 0xCB;
        if (b && where && where->empty()) *where = std::string(g.what) + (side ? " (band behind)" : " (band in front)");
        bad += b;
      }
    }
    *bad_out = bad;
    return 0;
  }
  int device = 0;
  int B = 0, in_h = 0, in_w = 0, in_c = 0;
  int opt_kind = 1;
  bool use_graph = true;
  std::vector<BlockRt> blocks;
  long long nparam = 0;                  // padded float count, multiple of 32
  // arena: params | grads | slot0 | slot1 | flags | counters   (one allocation => one IPC handle)
  uint8_t* arena = nullptr; size_t arena_bytes = 0;
  size_t off_params = 0, off_grads = 0, off_slot0 = 0, off_slot1 = 0, off_flags = 0, off_ctr = 0;
  // push all-reduce (allreduce.cuh): per-bucket flag tables + counters, and the receive area peers store gradients into
  size_t off_flags2 = 0, off_ctr2 = 0, off_recv = 0, recv_cap4 = 0, oneshot_cap4 = 0;
  unsigned long long* ar_trace = nullptr;      // optional globaltimer stamps of the push all-reduce (cib_engine_ar_trace)
  EngineNvls* nv = nullptr;                    // multicast-bound gradient buffer (multi-process NVLS path), see nvls_drv.cuh
  float *params = nullptr, *grads = nullptr, *slot0 = nullptr, *slot1 = nullptr;
  unsigned int* flags = nullptr; unsigned int* epoch_ctr = nullptr;
  HyperBlock* hb = nullptr;
  float* xin[2] = {nullptr, nullptr};    // batch inputs (double buffered for the host-fed path)
  float* yb[2] = {nullptr, nullptr};
  float* sw = nullptr;                   // sample weights [B]
  float* probs = nullptr;                // [B][O]
  int label_w = 1;
  // dataset
  const float* data_x = nullptr; const float* data_y = nullptr; int n_data = 0; int* perm = nullptr; int perm_cap = 0;
  bool has_perm = false;
  // streams / graphs
  cudaStream_t stream = nullptr, side = nullptr, copy_stream = nullptr;
  // Host-fed steps split the input copy into CIB_COPY_CHUNKS pieces on as many streams: one 6.3 MB pinned H2D copy runs at
  // 23 GB/s on this class of box (0.275 ms, longer than the training step), four concurrent quarter-size copies at 46 GB/s.
  cudaStream_t copy_more[7] = {};
  cudaEvent_t ev_copy_more[2][7] = {};
  cudaEvent_t ev_early = nullptr, ev_early_wg = nullptr, ev_early_done = nullptr;   // bucket-A update on its own stream
  cudaStream_t upd = nullptr;
  float* head_scratch = nullptr;       // per-CTA loss / accuracy partials + ticket of the multi-CTA head kernel
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_copy[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
  cudaGraphExec_t g_train = nullptr, g_eval = nullptr, g_train_host[2] = {nullptr, nullptr};
  bool graphs_dirty = true;
  long long host_calls = 0;
  int stream_priority = 0;
  int kcount = 0, train_kernels = 0, eval_kernels = 0, host_kernels = 0;
  // DP
  Comm* comm = nullptr; int comm_rank = 0; int ar_mode = 0;
  float* loss_pin = nullptr;

  int create(const CibBlock* bl, int nb, int batch, int h, int w, int c, long long np, int opt, int flags);
  void destroy();
  int enqueue_forward(cudaStream_t st, int k, bool training, bool gather, float* probs_out);
  int enqueue_backward(cudaStream_t st, int k, bool early_update = false);
  int enqueue_update(cudaStream_t st);
  int enqueue_update_range(cudaStream_t st, long long start, long long count, bool advance);
  bool bucket_one_shot(long long count) const;
  bool bucket_ll(long long count) const;
  bool use_early_update() const;
  int gather_slots();
  long long split_at() const;          // first parameter of the dense stack (0 = no split): everything from here on is bucket A
  int enqueue_step(cudaStream_t st, int k, bool training, bool gather);
  int build_graphs();
  int capture(cudaGraphExec_t* out, int k, bool training, bool gather);
};

struct Comm {
  int world = 1;
  bool ipc = false;
  // per rank: mapped arena base pointers as seen from the owning process of this Comm object
  std::vector<std::vector<uint8_t*>> bases;   // bases[viewer_rank][peer_rank]  (local: all identical)
  std::vector<void*> opened;                  // IPC mappings to close
};

// ------------------------------------------------------------------------------------------------------------
int Engine::create(const CibBlock* bl, int nb, int batch, int h, int w, int c, long long np, int opt, int flags_) {
  CIB_CUDA(cudaGetDevice(&device));
  B = batch; in_h = h; in_w = w; in_c = c; opt_kind = opt; nparam = np; use_graph = (flags_ & 1) != 0;
  if (nparam % 32) return fail("engine: nparam must be a multiple of 32 floats");
  blocks.resize(nb);
  for (int i = 0; i < nb; ++i) blocks[i].b = bl[i];
  if (nb < 2 || blocks[nb - 1].b.kind != 2 || blocks[nb - 2].b.kind != 1)
    return fail("engine: the device path needs ... -> hidden Dense -> output Dense");
  if (blocks[nb - 1].b.N > CIB_HEAD_MAXO) return fail("engine: output head wider than 16 units");

  // ---- arena
  const size_t pbytes = align_up((size_t)nparam * 4, 1024);
  off_params = 0; off_grads = pbytes; off_slot0 = 2 * pbytes; off_slot1 = 3 * pbytes;
  off_flags = 4 * pbytes;
  const size_t flag_bytes = align_up((size_t)2 * CIB_AR_MAX_CTAS * CIB_MAX_RANKS * 4, 1024);
  off_ctr = off_flags + flag_bytes;
  off_flags2 = off_ctr + 1024;
  const size_t flag2_bytes = (size_t)2 /*buckets*/ * 2 * CIB_PAR_MAX_CTAS * CIB_MAX_RANKS * 4;
  off_ctr2 = off_flags2 + flag2_bytes;
  off_recv = off_ctr2 + 1024;
  // (LL buckets: 4 x world x bucket, at most CIB_LL_MAX4 float4 each, two buckets)
  // receive area: two-shot buckets need (bucket + world) float4 in total, one-shot buckets 2 x world x bucket (allowed up to 4 MB per
  // bucket); the world size is not known yet, so this is sized for 8 ranks
  oneshot_cap4 = std::min<size_t>((size_t)nparam / 4, (size_t)(4u << 20) / 16);
  recv_cap4 = (size_t)nparam / 4 + 64 + 2 * CIB_MAX_RANKS * oneshot_cap4 + 2 * 4 * CIB_MAX_RANKS * std::min<size_t>((size_t)nparam / 4, CIB_LL_MAX4);
  arena_bytes = off_recv + recv_cap4 * 16;
  CIB_TRY(dev_alloc(&arena, arena_bytes, "arena"));
  CIB_CUDA(cudaMemset(arena, 0, arena_bytes));
  params = (float*)(arena + off_params); grads = (float*)(arena + off_grads);
  slot0 = (float*)(arena + off_slot0); slot1 = (float*)(arena + off_slot1);
  flags = (unsigned int*)(arena + off_flags); epoch_ctr = (unsigned int*)(arena + off_ctr);

  CIB_TRY(dev_alloc(&head_scratch, 1024, "head_scratch")); CIB_CUDA(cudaMemset(head_scratch, 0, 1024));
  CIB_TRY(dev_alloc(&hb, sizeof(HyperBlock), "hyper block"));
  HyperBlock h0; memset(&h0, 0, sizeof h0);
  h0.lr = 1e-3f; h0.beta1 = 0.9f; h0.beta2 = 0.999f; h0.eps = 1e-7f; h0.rho = 0.95f; h0.schedule_decay = 0.004f;
  h0.m_schedule[0] = h0.m_schedule[1] = 1.f; h0.seed = 1234u; h0.training = 1;
  CIB_CUDA(cudaMemcpy(hb, &h0, sizeof h0, cudaMemcpyHostToDevice));

  // ---- batch buffers
  const size_t xin_n = (size_t)B * h * w * c;
  label_w = blocks[nb - 1].b.N;
  for (int k = 0; k < 2; ++k) {
    CIB_TRY(dev_alloc(&xin[k], xin_n * 4, "batch x")); CIB_CUDA(cudaMemset(xin[k], 0, xin_n * 4));
    CIB_TRY(dev_alloc(&yb[k], (size_t)B * label_w * 4, "batch y")); CIB_CUDA(cudaMemset(yb[k], 0, (size_t)B * label_w * 4));
  }
  CIB_TRY(dev_alloc(&sw, (size_t)B * 4, "sample weights"));
  { std::vector<float> ones(B, 1.f); CIB_CUDA(cudaMemcpy(sw, ones.data(), (size_t)B * 4, cudaMemcpyHostToDevice)); }
  CIB_TRY(dev_alloc(&probs, (size_t)B * label_w * 4, "probs"));
  CIB_CUDA(cudaHostAlloc(&loss_pin, (size_t)CIB_LOSS_RING * 2 * sizeof(float), cudaHostAllocMapped | cudaHostAllocPortable));
  memset(loss_pin, 0, (size_t)CIB_LOSS_RING * 2 * sizeof(float));

  // ---- per-block buffers
  for (int i = 0; i < nb; ++i) {
    BlockRt& r = blocks[i]; const CibBlock& b = r.b;
    if (b.kind == 0) {
      r.Hs = b.pool ? b.Ho / 2 : b.Ho; r.Ws = b.pool ? b.Wo / 2 : b.Wo;
      r.act_n = (long long)B * r.Hs * r.Ws * b.Co;
      r.from_pooled = (i == 0 && b.path == 0 && b.ksize == 3 && (b.Ci == 1 || b.Ci == 3) && b.Co <= 256 && 256 % b.Co == 0);
      r.dz_n = r.from_pooled ? r.act_n : (long long)B * b.Ho * b.Wo * b.Co;
    } else if (b.kind == 1) {
      r.act_n = (long long)B * b.N; r.dz_n = r.act_n;
      CIB_TRY(dev_alloc(&r.z, r.act_n * 4, "dense z")); CIB_CUDA(cudaMemset(r.z, 0, r.act_n * 4));
    }
    // +256 B of slack: the 16-channel wgrad tensor map reads one pixel past the last row (see prepare_wgrad)
    if (r.act_n) { CIB_TRY(dev_alloc(&r.act, r.act_n * 4 + 256, "activation")); CIB_CUDA(cudaMemset(r.act, 0, r.act_n * 4 + 256)); }
    if (r.dz_n)  { CIB_TRY(dev_alloc(&r.dz, r.dz_n * 4, "dz"));   CIB_CUDA(cudaMemset(r.dz, 0, r.dz_n * 4)); }
  }
  int lo = 0, hi = 0;
  CIB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CIB_CUDA(cudaStreamCreateWithPriority(&stream, cudaStreamNonBlocking, lo));
  CIB_CUDA(cudaStreamCreateWithPriority(&side, cudaStreamNonBlocking, lo));
  CIB_CUDA(cudaStreamCreateWithPriority(&copy_stream, cudaStreamNonBlocking, lo));
  for (int c = 0; c < 7; ++c) CIB_CUDA(cudaStreamCreateWithPriority(&copy_more[c], cudaStreamNonBlocking, lo));
  for (int k = 0; k < 2; ++k) for (int c = 0; c < 7; ++c) CIB_CUDA(cudaEventCreateWithFlags(&ev_copy_more[k][c], cudaEventDisableTiming));
  CIB_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
  CIB_CUDA(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
  CIB_CUDA(cudaEventCreateWithFlags(&ev_early, cudaEventDisableTiming));
  CIB_CUDA(cudaEventCreateWithFlags(&ev_early_wg, cudaEventDisableTiming));
  CIB_CUDA(cudaEventCreateWithFlags(&ev_early_done, cudaEventDisableTiming));
  // the early bucket's few small CTAs must get onto the SMs ahead of the queued convolution CTAs: greatest priority
  CIB_CUDA(cudaStreamCreateWithPriority(&upd, cudaStreamNonBlocking, hi));
  for (int k = 0; k < 2; ++k) {
    CIB_CUDA(cudaEventCreateWithFlags(&ev_copy[k], cudaEventDisableTiming));
    CIB_CUDA(cudaEventCreateWithFlags(&ev_done[k], cudaEventDisableTiming));
  }
  return 0;
}

void Engine::destroy() {
  if (stream) cudaStreamSynchronize(stream);
  for (auto g : {g_train, g_eval, g_train_host[0], g_train_host[1]}) if (g) cudaGraphExecDestroy(g);
  for (auto& r : blocks) { dev_free(r.act); dev_free(r.dz); dev_free(r.z); }
  for (int k = 0; k < 2; ++k) { dev_free(xin[k]); dev_free(yb[k]); if (ev_copy[k]) cudaEventDestroy(ev_copy[k]); if (ev_done[k]) cudaEventDestroy(ev_done[k]); }
  dev_free(head_scratch);
  if (nv) { engine_nvls_destroy(nv); nv = nullptr; }
  dev_free(sw); dev_free(probs); dev_free(hb); cudaFree(perm); dev_free(arena);
  if (loss_pin) cudaFreeHost(loss_pin);
  if (ev_fork) cudaEventDestroy(ev_fork); if (ev_join) cudaEventDestroy(ev_join); if (ev_early) cudaEventDestroy(ev_early); if (ev_early_wg) cudaEventDestroy(ev_early_wg);
  if (ev_early_done) cudaEventDestroy(ev_early_done); if (upd) cudaStreamDestroy(upd);
  if (stream) cudaStreamDestroy(stream); if (side) cudaStreamDestroy(side); if (copy_stream) cudaStreamDestroy(copy_stream);
  for (int c = 0; c < 7; ++c) if (copy_more[c]) cudaStreamDestroy(copy_more[c]);
  for (int k = 0; k < 2; ++k) for (int c = 0; c < 7; ++c) if (ev_copy_more[k][c]) cudaEventDestroy(ev_copy_more[k][c]);
}

}  // namespace cib
