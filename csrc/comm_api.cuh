// Communicators over peer-mapped arenas (same-process peer access or CUDA IPC) and stand-alone symmetric buffers.
// Replaces hvd.init() / MPI communicator bootstrap (reference train_rpv.py:37, rpv.py:63-65).
#pragma once
#include "engine_graph.cuh"

using namespace cib;

namespace cib {

struct SymBuf {
  int device = 0;
  long long nfloats = 0;
  // arena: data | scratch | flags | ctr
  uint8_t* arena = nullptr; size_t bytes = 0, off_scratch = 0, off_flags = 0, off_ctr = 0;
  int rank = 0, world = 1;
  uint8_t* peers[CIB_MAX_RANKS] = {nullptr};
  std::vector<void*> opened;
};

static int enable_peer(int dev, int peer) {
  if (dev == peer) return 0;
  int can = 0;
  CIB_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer));
  if (!can) return fail("peer access not possible between devices " + std::to_string(dev) + " and " + std::to_string(peer));
  CIB_CUDA(cudaSetDevice(dev));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(std::string("cudaDeviceEnablePeerAccess -> ") + cudaGetErrorString(e));
  cudaGetLastError();
  return 0;
}

}  // namespace cib

extern "C" {

int cib_comm_create_local(cib_comm_t* out, cib_engine_t* engines, int world) {
  if (world < 1 || world > CIB_MAX_RANKS) return fail("comm: world must be 1..8");
  Comm* c = new Comm(); c->world = world; c->ipc = false;
  c->bases.assign(world, std::vector<uint8_t*>(world, nullptr));
  int cur = 0; cudaGetDevice(&cur);
  for (int r = 0; r < world; ++r)
    for (int q = 0; q < world; ++q) {
      Engine* er = reinterpret_cast<Engine*>(engines[r]); Engine* eq = reinterpret_cast<Engine*>(engines[q]);
      if (enable_peer(er->device, eq->device)) { delete c; cudaSetDevice(cur); return 1; }
      c->bases[r][q] = eq->arena;
    }
  cudaSetDevice(cur);
  *out = c;
  return 0;
}

int cib_engine_ipc_handle(cib_engine_t e_, unsigned char handle[64]) {
  Engine* e = reinterpret_cast<Engine*>(e_);
  cudaIpcMemHandle_t h;
  CIB_CUDA(cudaIpcGetMemHandle(&h, e->arena));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle, &h, 64);
  return 0;
}

int cib_comm_create_ipc(cib_comm_t* out, cib_engine_t e_, int rank, int world, const unsigned char* handles) {
  if (world < 1 || world > CIB_MAX_RANKS) return fail("comm: world must be 1..8");
  Engine* e = reinterpret_cast<Engine*>(e_);
  Comm* c = new Comm(); c->world = world; c->ipc = true;
  c->bases.assign(1, std::vector<uint8_t*>(world, nullptr));
  for (int q = 0; q < world; ++q) {
    if (q == rank) { c->bases[0][q] = e->arena; continue; }
    cudaIpcMemHandle_t h; memcpy(&h, handles + 64 * q, 64);
    void* p = nullptr;
    cudaError_t err = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (err != cudaSuccess) { delete c; return fail(std::string("cudaIpcOpenMemHandle(rank ") + std::to_string(q) + ") -> " + cudaGetErrorString(err)); }
    // an IPC mapping starts at the peer's ALLOCATION; with canary bands (CIB_DEBUG_BOUNDS) the arena begins one band later
    c->bases[0][q] = (uint8_t*)p + (g_debug_bounds ? CIB_CANARY : 0); c->opened.push_back(p);
  }
  *out = c;
  return 0;
}

int cib_comm_destroy(cib_comm_t c_) {
  Comm* c = reinterpret_cast<Comm*>(c_);
  if (!c) return 0;
  for (void* p : c->opened) cudaIpcCloseMemHandle(p);
  delete c;
  return 0;
}

int cib_engine_attach_comm(cib_engine_t e_, cib_comm_t c_, int rank, int mode) {
  Engine* e = reinterpret_cast<Engine*>(e_);
  e->comm = reinterpret_cast<Comm*>(c_); e->comm_rank = rank; e->ar_mode = mode; e->graphs_dirty = true;
  return 0;
}

// Broadcast params + both optimizer slots from root: the three buffers are contiguous in the arena apart from the
// gradient buffer, which is zero everywhere, so one pull of [params .. slot1] does it.
int cib_engine_broadcast_state(cib_engine_t e_, int root) {
  Engine* e = reinterpret_cast<Engine*>(e_);
  if (!e->comm || e->comm->world < 2) return 0;
  CIB_CUDA(cudaSetDevice(e->device));
  AllreduceParams p{};
  CIB_TRY(fill_commdev(&p.comm, e->comm, e->comm_rank, e->off_params, e->off_params, e->off_flags));
  p.n4 = (int)((e->off_flags - e->off_params) / 16);
  p.epoch_ctr = e->epoch_ctr; p.ticket = e->epoch_ctr + 1;
  broadcast_kernel<<<ar_grid_for(p.n4), 512, 0, e->stream>>>(p, root);
  CIB_CUDA(cudaGetLastError());
  return 0;
}

// Optimizer moments are sharded by the two-shot update (each rank updates its own slice): reassemble them on every rank.
// Collective -- all ranks call it at the same point (end of an epoch, before a checkpoint / broadcast / engine re-creation).
int cib_engine_gather_slots(cib_engine_t e_) {
  Engine* e = reinterpret_cast<Engine*>(e_);
  if (!e->comm || e->comm->world < 2) return 0;
  CIB_CUDA(cudaSetDevice(e->device));
  return e->gather_slots();
}

// ---- NVLS across processes (see nvls_drv.cuh for the protocol; parallel/dist.py drives it with host barriers in between) ----------
// step 0 (all): size;  step 1: rank 0 creates + exports (*fd_out), the others import fd_in;  step 2 (all): add the own device;
// step 3 (all, after a barrier): bind + map, the engine's gradient buffer moves into the multicast-bound memory.
int cib_engine_nvls_step(cib_engine_t e_, int step, int world, int fd_in, int* fd_out) {
  Engine* e = reinterpret_cast<Engine*>(e_);
  CIB_CUDA(cudaSetDevice(e->device));
  std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
  if (step == 0) {
    if (e->nv) { engine_nvls_destroy(e->nv); e->nv = nullptr; }
    e->nv = new EngineNvls();
    int rc = engine_nvls_size(e->nv, e->device, world, (size_t)e->nparam * 4);
    if (rc) { engine_nvls_destroy(e->nv); e->nv = nullptr; }
    return rc;
  }
  if (!e->nv) return fail("nvls: step 0 has not run");
  if (step == 1) return fd_out ? engine_nvls_create(e->nv, fd_out) : engine_nvls_import(e->nv, fd_in);
  if (step == 2) return engine_nvls_add_device(e->nv);
  if (step == 3) {
    CIB_CUDA(cudaStreamSynchronize(e->stream));
    CIB_TRY(engine_nvls_bind_map(e->nv));
    e->grads = (float*)e->nv->uc;          // the wgrad kernels now accumulate into the multicast-bound buffer
    e->graphs_dirty = true;
    return 0;
  }
  if (step == 9) { engine_nvls_destroy(e->nv); e->nv = nullptr; e->grads = (float*)(e->arena + e->off_grads); e->graphs_dirty = true; return 0; }
  return fail("nvls: unknown step");
}
int cib_engine_nvls_active(cib_engine_t e_) { Engine* e = reinterpret_cast<Engine*>(e_); return e->nv && e->nv->ready ? 1 : 0; }

// In-kernel timeline of the push all-reduce (bucket 0 / bucket 1, CTA 0): out[b*8 + {0 start, 1 flags seen after push, 2 reduced +
// updated (+ parameters pushed), 3 parameter flags seen}] in nanoseconds of %globaltimer, plus step-level stamps written by one-thread
// marker kernels: out[16] step start, [17] early bucket eligible (dense wgrad + dgrad done), [18] early bucket done, [19] backward
// joined (last bucket eligible), [20] last bucket done.  enable: 1 allocates and switches it on for graphs captured from now on
// (the engine re-captures, with the 5 marker kernels added), 0 only reads.  out32 (may be NULL) receives 32 values.
int cib_engine_ar_trace(cib_engine_t e_, int enable, unsigned long long* out16) {
  Engine* e = reinterpret_cast<Engine*>(e_);
  CIB_CUDA(cudaSetDevice(e->device));
  if (enable && !e->ar_trace) {
    std::lock_guard<std::recursive_mutex> lock(g_exclusive_mu);
    CIB_CUDA(cudaMalloc(&e->ar_trace, 32 * 8)); CIB_CUDA(cudaMemset(e->ar_trace, 0, 32 * 8));
    e->graphs_dirty = true;
  }
  if (out16) {
    if (!e->ar_trace) return fail("ar_trace: not enabled");
    CIB_CUDA(cudaStreamSynchronize(e->stream));
    CIB_CUDA(cudaMemcpy(out16, e->ar_trace, 32 * 8, cudaMemcpyDeviceToHost));
  }
  return 0;
}

// --------------------------------------------------------------------------------------------------------
int cib_symbuf_create(cib_symbuf_t* out, int device, long long nfloats) {
  SymBuf* s = new SymBuf(); s->device = device; s->nfloats = nfloats;
  CIB_CUDA(cudaSetDevice(device));
  const size_t dbytes = align_up((size_t)nfloats * 4, 1024);
  const size_t sbytes = std::min<size_t>(dbytes, 1 << 20);        // one-shot staging is only used for small messages
  s->off_scratch = dbytes; s->off_flags = dbytes + sbytes;
  s->off_ctr = s->off_flags + align_up((size_t)2 * CIB_AR_MAX_CTAS * CIB_MAX_RANKS * 4, 1024);
  s->bytes = s->off_ctr + 1024;
  CIB_CUDA(cudaMalloc(&s->arena, s->bytes));
  CIB_CUDA(cudaMemset(s->arena, 0, s->bytes));
  s->peers[0] = s->arena;
  *out = s;
  return 0;
}
int cib_symbuf_ipc_handle(cib_symbuf_t s_, unsigned char handle[64]) {
  SymBuf* s = reinterpret_cast<SymBuf*>(s_);
  cudaIpcMemHandle_t h;
  CIB_CUDA(cudaIpcGetMemHandle(&h, s->arena));
  memcpy(handle, &h, 64);
  return 0;
}
int cib_symbuf_attach_ipc(cib_symbuf_t s_, int rank, int world, const unsigned char* handles) {
  SymBuf* s = reinterpret_cast<SymBuf*>(s_);
  s->rank = rank; s->world = world;
  for (int q = 0; q < world; ++q) {
    if (q == rank) { s->peers[q] = s->arena; continue; }
    cudaIpcMemHandle_t h; memcpy(&h, handles + 64 * q, 64);
    void* p = nullptr;
    CIB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    s->peers[q] = (uint8_t*)p; s->opened.push_back(p);
  }
  return 0;
}
int cib_symbuf_attach_local(cib_symbuf_t* bufs, int world) {
  int cur = 0; cudaGetDevice(&cur);
  for (int r = 0; r < world; ++r) {
    SymBuf* s = reinterpret_cast<SymBuf*>(bufs[r]);
    s->rank = r; s->world = world;
    for (int q = 0; q < world; ++q) {
      SymBuf* o = reinterpret_cast<SymBuf*>(bufs[q]);
      if (enable_peer(s->device, o->device)) { cudaSetDevice(cur); return 1; }
      s->peers[q] = o->arena;
    }
  }
  cudaSetDevice(cur);
  return 0;
}
int cib_symbuf_ptr(cib_symbuf_t s_, void** dev_ptr) { *dev_ptr = reinterpret_cast<SymBuf*>(s_)->arena; return 0; }

static int symbuf_params(SymBuf* s, long long nfloats, AllreduceParams* p) {
  if (nfloats > s->nfloats || nfloats % 4) return fail("symbuf: size must be a multiple of 4 floats within the buffer");
  memset(p, 0, sizeof *p);
  p->comm.rank = s->rank; p->comm.world = s->world; p->comm.timeout_ns = g_peer_timeout_ns;
  for (int q = 0; q < s->world; ++q) {
    p->comm.grads[q] = (float*)s->peers[q];
    p->comm.params[q] = (float*)s->peers[q];
    p->comm.flags[q] = (unsigned int*)(s->peers[q] + s->off_flags);
  }
  p->n4 = (int)(nfloats / 4);
  p->epoch_ctr = (unsigned int*)(s->arena + s->off_ctr); p->ticket = p->epoch_ctr + 1;
  p->scratch = (float*)(s->arena + s->off_scratch);
  return 0;
}
int cib_symbuf_allreduce(cib_symbuf_t s_, long long nfloats, int average, int mode, void* stream) {
  SymBuf* s = reinterpret_cast<SymBuf*>(s_);
  AllreduceParams p;
  CIB_CUDA(cudaSetDevice(s->device));
  CIB_TRY(symbuf_params(s, nfloats, &p));
  p.average = average;
  const long long bytes = nfloats * 4;
  p.one_shot = mode == 1 ? 1 : mode == 2 ? 0 : (bytes <= 256 * 1024 ? 1 : 0);
  if (p.one_shot && bytes > (1 << 20)) p.one_shot = 0;
  allreduce_kernel<<<ar_grid_for(p.n4), 512, 0, (cudaStream_t)stream>>>(p);
  CIB_CUDA(cudaGetLastError());
  return 0;
}
int cib_symbuf_broadcast(cib_symbuf_t s_, long long nfloats, int root, void* stream) {
  SymBuf* s = reinterpret_cast<SymBuf*>(s_);
  AllreduceParams p;
  CIB_CUDA(cudaSetDevice(s->device));
  CIB_TRY(symbuf_params(s, nfloats, &p));
  broadcast_kernel<<<ar_grid_for(p.n4), 512, 0, (cudaStream_t)stream>>>(p, root);
  CIB_CUDA(cudaGetLastError());
  return 0;
}
int cib_symbuf_destroy(cib_symbuf_t s_) {
  SymBuf* s = reinterpret_cast<SymBuf*>(s_);
  if (!s) return 0;
  for (void* p : s->opened) cudaIpcCloseMemHandle(p);
  cudaFree(s->arena);
  delete s;
  return 0;
}

}  // extern "C"
