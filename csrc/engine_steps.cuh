// Kernel sequencing of one step (forward, backward, update) and CUDA-graph capture.
// Layer chains served: mnist.py:48-56 (conv, conv, pool, dropout, dense, dropout, softmax), rpv.py:42-59 (conv+pool blocks,
// dropout, dense blocks, sigmoid), Train_rpv.ipynb:204-218 (strided legacy model); ops K1-K12 of SURVEY 2.7.1.
#pragma once
#include "engine.cuh"

namespace cib {

static inline int grid_for(long long total, int per_block = 256, int cap_mult = 8) {
  long long b = (total + per_block - 1) / per_block;
  if (b < 1) b = 1;
  return (int)std::min<long long>(b, (long long)g_num_sms * cap_mult);
}

int Engine::enqueue_forward(cudaStream_t st, int k, bool training, bool gather, float* probs_out) {
  const int nb = (int)blocks.size();
  if (gather) {
    GatherParams g{};
    g.xb = xin[k]; g.yb = yb[k]; g.sw = sw;
    g.row_f4 = in_h * in_w * in_c / 4; g.label_w = label_w; g.B = B; g.hb = hb;
    if ((in_h * in_w * in_c) % 4) return fail("engine: sample size must be a multiple of 4 floats");
    dim3 grid(std::max(1, std::min(8, g.row_f4 / 256)), B);
    gather_batch_kernel<<<grid, 256, 0, st>>>(g); ++kcount;
    CIB_CUDA(cudaGetLastError());
  }
  // programmatic dependent launch: allowed for a kernel whose predecessor in `st` is a kernel of this same enqueue sequence (the gather
  // kernel, or the previous block); the first kernel of a host-fed step follows copies / event waits and is launched the ordinary way
  bool pdl_ok = gather;
  for (int i = 0; i < nb; ++i, pdl_ok = true) {
    BlockRt& r = blocks[i]; const CibBlock& b = r.b;
    const float* src = i == 0 ? xin[k] : blocks[i - 1].act;
    const float drop = training ? b.drop_rate : 0.f;
    if (b.kind == 0) {
      if (b.path == 1) {
        ConvEpilogue e;
        e.epi = CONV_EPI_FWD; e.bias = params + b.b_off; e.out = r.act; e.pool = b.pool; e.drop_rate = drop; e.layer_id = i; e.hb = hb;
        t_pdl = pdl_ok; CIB_TRY(run_conv(true, src, B, b.Hi, b.Wi, b.Ci, params + b.w_off, b.Ci, b.Co, b.Ho, b.Wo, b.pad, e, st, b.stride)); t_pdl = false; ++kcount;
      } else {
        if (i == 0 && conv1_tc_ok(b)) {
          Conv1TcParams p{};
          p.x = src; p.w = params + b.w_off; p.bias = params + b.b_off; p.out = r.act;
          p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Ho = b.Ho; p.Wo = b.Wo; p.pad = b.pad; p.pool = b.pool;
          p.drop_rate = drop; p.layer_id = i; p.hb = hb;
          t_pdl = pdl_ok; CIB_TRY(launch_conv1_tc(p, false, st)); t_pdl = false; ++kcount;
        } else if (first_tiled_ok(b)) {
          ConvFirstParams p{};
          p.x = src; p.w = params + b.w_off; p.bias = params + b.b_off; p.out = r.act;
          p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Ho = b.Ho; p.Wo = b.Wo; p.pad = b.pad;
          p.drop_rate = drop; p.layer_id = i; p.hb = hb;
          t_pdl = pdl_ok; CIB_TRY(launch_conv_first(p, false, st)); t_pdl = false; ++kcount;
        } else {
          ConvDirectParams p{};
          p.x = src; p.w = params + b.w_off; p.bias = params + b.b_off; p.out = r.act;
          p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Ho = b.Ho; p.Wo = b.Wo;
          p.KH = p.KW = b.ksize; p.stride = b.stride; p.pad = b.pad; p.pool = b.pool;
          p.drop_rate = drop; p.layer_id = i; p.hb = hb;
          CIB_TRY(launch_conv_direct(p, st)); ++kcount;
        }
      }
    } else if (b.kind == 1) {
      if (b.path == 1) {
        GemmLaunch L;
        const int nk = (b.K + 31) / 32;
        const int splits = std::max(1, std::min(nk / 4, g_num_sms));
        CIB_TRY(prepare_gemm(&L, src, b.K, params + b.w_off, b.N, r.z, b.N, B, b.N, b.K, 0, 1, splits, GEMM_EPI_REDADD, 0));
        t_pdl = pdl_ok; CIB_TRY(launch_gemm(L, st)); t_pdl = false; ++kcount;
      } else {
        DenseGenericParams p{};
        p.a = src; p.w = params + b.w_off; p.z = r.z; p.M = B; p.N = b.N; p.K = b.K;
        dense_fwd_generic_kernel<<<grid_for((long long)B * b.N), 256, 0, st>>>(p); ++kcount;
        CIB_CUDA(cudaGetLastError());
      }
      // the last hidden layer's bias + ReLU + dropout is applied by its consumer, the multi-CTA loss head (see HeadParams::z)
      const bool act_in_head = g_fuse_dense_act && g_head_version == 2 && i == nb - 2;
      if (!act_in_head) {
        t_pdl = pdl_ok;
        CIB_CUDA(launch_k(bias_relu_dropout_kernel, dim3(grid_for((long long)B * b.N)), dim3(256), 0, st, r.z, (const float*)(params + b.b_off), r.act, B, b.N, drop,
                          (unsigned int)i, (const HyperBlock*)hb)); ++kcount;
        CIB_CUDA(cudaGetLastError());
      }
    } else {
      const BlockRt& h = blocks[i - 1];
      HeadParams p{};
      p.a = h.act; p.w = params + b.w_off; p.bias = params + b.b_off; p.y = yb[k]; p.sw = sw;
      p.probs = probs_out ? probs_out : probs;
      p.dW = grads + b.w_off; p.db = grads + b.b_off;
      p.dz_hidden = h.dz; p.db_hidden = grads + h.b.b_off;
      p.hb = hb; p.B = B; p.F = b.K; p.O = b.N; p.kind = b.head_kind; p.train = training ? 1 : 0;
      p.advance_cursor = gather ? 1 : 0;
      if (!gather && training) {                       // host-fed step: result goes to the pinned host ring as well
        float* dev_view = nullptr;
        if (cudaHostGetDevicePointer((void**)&dev_view, loss_pin, 0) == cudaSuccess) { p.loss_ring = dev_view; p.ring_mask = CIB_LOSS_RING - 1; }
      }
      p.hidden_scale = training && h.b.drop_rate > 0.f ? 1.f / (1.f - h.b.drop_rate) : 1.f;
      if (b.K > 1024) return fail("head: hidden layer wider than 1024 units");
      size_t smem = ((((size_t)B * b.N + 3) & ~(size_t)3) + 128) * 4;
      const size_t a_bytes = (size_t)B * b.K * 4;
      const size_t part_bytes = (size_t)(1024 / b.K) * b.K * (b.N + 1) * 4;      // [G][F][O + 1] partial sums
      p.a_in_smem = (b.K % 4 == 0 && smem + a_bytes + part_bytes <= 200 * 1024) ? 1 : 0;
      if (p.a_in_smem) smem += a_bytes;
      smem += part_bytes;
      if (g_head_version == 2) {
        if (g_fuse_dense_act) {
          p.z = h.z; p.bias_hidden = params + h.b.b_off; p.act_out = h.act; p.drop_hidden = training ? h.b.drop_rate : 0.f; p.layer_hidden = (unsigned int)(i - 1);
        }
        // multi-CTA head: ~8 samples per CTA
        const int nc = std::max(1, std::min(64, (B + g_head_spc - 1) / g_head_spc));     // g_head_spc samples per CTA
        const int S = (B + nc - 1) / nc;
        const size_t smem2 = ((size_t)S * b.K + (size_t)b.K * b.N + (((size_t)S * b.N + 3) & ~(size_t)3) + 64) * 4;
        t_pdl = pdl_ok;
        if (b.N == 1) CIB_CUDA(launch_k(head_mc_kernel<1>, dim3(nc), dim3(256), smem2, st, p, head_scratch));
        else CIB_CUDA(launch_k(head_mc_kernel<16>, dim3(nc), dim3(256), smem2, st, p, head_scratch));
      } else {
        t_pdl = pdl_ok;
        if (b.N == 1) CIB_CUDA(launch_k(head_kernel<1>, dim3(1), dim3(1024), smem, st, p)); else CIB_CUDA(launch_k(head_kernel<16>, dim3(1), dim3(1024), smem, st, p));
      }
      ++kcount;
      CIB_CUDA(cudaGetLastError());
    }
  }
  return 0;
}

// Backward of blocks nb-2 .. 0 (the head's own backward ran inside head_kernel).  wgrad kernels go to the side
// stream so that, inside the captured graph, each layer's wgrad and dgrad are independent branches.
int Engine::enqueue_backward(cudaStream_t st, int k, bool early_update) {
  const int nb = (int)blocks.size();
  int first_dense = -1; bool early_pending = false;
  for (int i = 0; i < nb; ++i) if (blocks[i].b.kind != 0) { first_dense = i; break; }
  for (int j = nb - 2; j >= 0; --j) {
    BlockRt& r = blocks[j]; const CibBlock& b = r.b;
    const float* src = j == 0 ? xin[k] : blocks[j - 1].act;
    // dz of block j is complete on `st` here: let the side stream see it.  Convolution wgrads may instead stay on the main stream
    // (g_serial_bwd), chained to their neighbours by programmatic dependent launch
    const bool serial = g_serial_bwd && b.kind == 0;
    cudaStream_t ws = serial ? st : side;
    if (!serial) {
      CIB_CUDA(cudaEventRecord(ev_fork, st));
      CIB_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
    }
    // ------------------------------ wgrad (side stream) ------------------------------
    if (b.kind == 1) {
      if (b.path == 1) {
        GemmLaunch L;
        CIB_TRY(prepare_gemm(&L, src, b.K, r.dz, b.N, grads + b.w_off, b.N, b.K, b.N, B, 1, 1, 1, GEMM_EPI_STORE, 0));
        CIB_TRY(launch_gemm(L, side)); ++kcount;
      } else {
        DenseGenericParams p{};
        p.a = src; p.dz = r.dz; p.dW = grads + b.w_off; p.M = B; p.N = b.N; p.K = b.K;
        dense_wgrad_generic_kernel<<<grid_for((long long)b.K * b.N), 256, 0, side>>>(p); ++kcount;
        CIB_CUDA(cudaGetLastError());
      }
      if (j != nb - 2) {   // the last hidden layer's bias gradient came from the head kernel
        colsum_add_kernel<<<(b.N + 127) / 128, 128, 0, side>>>(r.dz, grads + b.b_off, B, b.N); ++kcount;
        CIB_CUDA(cudaGetLastError());
      }
    } else {
      if (j == 0 && r.from_pooled && conv1_tc_ok(b)) {
        Conv1TcParams p{};
        p.x = src; p.g = r.dz; p.act = r.act; p.dW = grads + b.w_off; p.db = grads + b.b_off;
        p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Ho = b.Ho; p.Wo = b.Wo; p.pad = b.pad; p.pool = b.pool;
        CIB_TRY(launch_conv1_tc(p, true, ws)); ++kcount;
      } else if (r.from_pooled && first_tiled_ok(b)) {
        ConvFirstParams p{};
        p.x = src; p.g = r.dz; p.act = r.act; p.dW = grads + b.w_off; p.db = grads + b.b_off;
        p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Ho = b.Ho; p.Wo = b.Wo; p.pad = b.pad;
        t_pdl = serial; CIB_TRY(launch_conv_first(p, true, ws)); t_pdl = false; ++kcount;
      } else if (r.from_pooled) {
        WgradFirstParams p{};
        p.x = src; p.g = r.dz; p.act = r.act; p.dW = grads + b.w_off; p.db = grads + b.b_off;
        p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Hs = r.Hs; p.Ws = r.Ws;
        p.pool = b.pool; p.pad = b.pad; p.stride = b.stride;
        CIB_TRY(launch_wgrad_first(p, ws)); ++kcount;
      } else if (b.path == 1) {
        t_pdl = serial; CIB_TRY(run_wgrad(src, r.dz, grads + b.w_off, grads + b.b_off, B, b.Hi, b.Wi, b.Ci, b.Ho, b.Wo, b.Co, b.pad, ws, b.stride)); t_pdl = false; ++kcount;
      } else {
        ConvGenericBwdParams p{};
        p.x = src; p.dz = r.dz; p.dW = grads + b.w_off; p.db = grads + b.b_off;
        p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Ho = b.Ho; p.Wo = b.Wo;
        p.KH = p.KW = b.ksize; p.stride = b.stride; p.pad = b.pad;
        dim3 grid(b.ksize * b.ksize * b.Ci + 1, 32);
        conv_wgrad_generic_kernel<<<grid, 256, 0, ws>>>(p); ++kcount;
        CIB_CUDA(cudaGetLastError());
      }
    }
    // ------------------------------ dgrad (main stream) ------------------------------
    if (j == 0) break;
    BlockRt& pr = blocks[j - 1]; const CibBlock& pb = pr.b;
    const float scale = pb.drop_rate > 0.f ? 1.f / (1.f - pb.drop_rate) : 1.f;
    const bool unpool = pb.kind == 0 && pb.pool && !pr.from_pooled;
    if (b.kind == 1) {
      if (b.path == 1 && (!unpool || (pb.Co <= 256 && pb.Co % 16 == 0))) {
        GemmLaunch L;
        // N tile of the un-pool epilogue: a divisor of the channel count; narrow enough that the grid fills the SMs (g_dgrad_bn overrides)
        int BN = 0;
        if (unpool) {
          BN = pb.Co;
          const int m_tiles = (B + 127) / 128;
          while (BN % 32 == 0 && BN > 16 && (long long)m_tiles * (b.K / BN) < g_num_sms) BN /= 2;
          if (g_dgrad_bn > 0 && pb.Co % g_dgrad_bn == 0) BN = g_dgrad_bn;
        }
        CIB_TRY(prepare_gemm(&L, r.dz, b.N, params + b.w_off, b.N, nullptr, b.K, B, b.K, b.N, 0, 0, 1,
                             unpool ? GEMM_EPI_UNPOOL : GEMM_EPI_MASK, BN));
        L.p.act_prev = pr.act; L.p.dz_out = pr.dz; L.p.scale = scale;
        if (unpool) { L.p.Hp = pr.Hs; L.p.Wp = pr.Ws; L.p.Cp = pb.Co; L.p.Hf = pb.Ho; L.p.Wf = pb.Wo; }
        t_pdl = true; CIB_TRY(launch_gemm(L, st)); t_pdl = false; ++kcount;
      } else {
        DenseGenericParams p{};
        p.dz = r.dz; p.w = params + b.w_off; p.act_prev = pr.act; p.dx_out = pr.dz; p.scale = scale;
        p.epi = unpool ? CONV_EPI_DGRAD_UNPOOL : CONV_EPI_DGRAD_MASK;
        if (pb.kind == 0) { p.Hp = pr.Hs; p.Wp = pr.Ws; p.Cp = pb.Co; p.Hf = pb.Ho; p.Wf = pb.Wo; }
        p.M = B; p.N = b.N; p.K = b.K;
        dense_dgrad_generic_kernel<<<grid_for((long long)B * b.K), 256, 0, st>>>(p); ++kcount;
        CIB_CUDA(cudaGetLastError());
      }
    } else {
      if (b.path == 1) {
        ConvEpilogue e;
        e.epi = unpool ? CONV_EPI_DGRAD_UNPOOL : CONV_EPI_DGRAD_MASK;
        e.act_prev = pr.act; e.dz_out = pr.dz; e.Hf = pb.Ho; e.Wf = pb.Wo; e.scale = scale; e.hb = hb;
        t_pdl = true; CIB_TRY(run_conv(false, r.dz, B, b.Ho, b.Wo, b.Co, params + b.w_off, b.Ci, b.Co, b.Hi, b.Wi, b.pad, e, st, b.stride)); t_pdl = false;
        kcount += b.stride == 2 ? 4 : 1;        // a strided dgrad is four launches, one per parity class
      } else {
        ConvGenericBwdParams p{};
        p.dz = r.dz; p.w = params + b.w_off; p.act_prev = pr.act; p.dx_out = pr.dz; p.scale = scale;
        p.epi = unpool ? CONV_EPI_DGRAD_UNPOOL : CONV_EPI_DGRAD_MASK; p.Hf = pb.Ho; p.Wf = pb.Wo;
        p.n_img = B; p.Hi = b.Hi; p.Wi = b.Wi; p.CI = b.Ci; p.CO = b.Co; p.Ho = b.Ho; p.Wo = b.Wo;
        p.KH = p.KW = b.ksize; p.stride = b.stride; p.pad = b.pad;
        conv_dgrad_generic_kernel<<<grid_for((long long)B * b.Hi * b.Wi * b.Ci), 256, 0, st>>>(p); ++kcount;
        CIB_CUDA(cudaGetLastError());
      }
    }
    if (early_update && j == first_dense) {
      // every gradient of the dense stack is complete (head kernel, the wgrads enqueued on `side` so far) and the dense
      // weights have been read for the last time by the dgrad just enqueued on `st`: update bucket A now on its own stream,
      // concurrently with the convolution backward (and without delaying the conv wgrads queued on `side`)
      const long long split = split_at();
      CIB_CUDA(cudaEventRecord(ev_early, st));
      CIB_CUDA(cudaEventRecord(ev_early_wg, side));
      CIB_CUDA(cudaStreamWaitEvent(upd, ev_early, 0));
      CIB_CUDA(cudaStreamWaitEvent(upd, ev_early_wg, 0));
      if (ar_trace) { stamp_kernel<<<1, 1, 0, upd>>>(ar_trace + 17); }
      CIB_TRY(enqueue_update_range(upd, split, nparam - split, false));
      if (ar_trace) { stamp_kernel<<<1, 1, 0, upd>>>(ar_trace + 18); }
      CIB_CUDA(cudaEventRecord(ev_early_done, upd));
      early_pending = true;
    }
  }
  // join the side stream (and the early update) back
  CIB_CUDA(cudaEventRecord(ev_join, side));
  CIB_CUDA(cudaStreamWaitEvent(st, ev_join, 0));
  if (early_pending) CIB_CUDA(cudaStreamWaitEvent(st, ev_early_done, 0));
  return 0;
}

}  // namespace cib
