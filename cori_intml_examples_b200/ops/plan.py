"""Planner: layer list -> fused device blocks (``CibBlock``) + padded flat-parameter layout.

The device engine works on a *program* of fused blocks:

    conv block   = Conv2D(relu) [+ MaxPooling2D(2,2)] [+ Dropout]
    hidden block = Dense(relu) [+ Dropout]
    head block   = Dense(sigmoid | softmax) + cross-entropy + accuracy

Channels / features are zero-padded in device memory to what tcgen05 + TMA need (16 for K-major-only tensors,
multiples of 32 otherwise); the logical (Keras) shapes are kept for checkpoints, ``get_weights`` and ``summary``.
Padding is self-preserving: padded weights start at zero, always see zero gradients, and every optimizer maps
(param = 0, grad = 0, slots = 0) to itself.

This module is pure Python/NumPy so the plan (shapes, offsets, packing) is unit-tested on CPU.
Reference topologies: mnist.py:48-56, rpv.py:42-59, Train_rpv.ipynb:209-211 (strided legacy model).
"""
import os

import numpy as np

from ..nn import layers as L

KIND_CONV, KIND_HIDDEN, KIND_HEAD = 0, 1, 2


class PlanError(ValueError):
    """The layer chain cannot run on the fused device path (the CPU oracle still can)."""


def _ceil(v, m):
    return (v + m - 1) // m * m


def pad_channels(c, allow16=True):
    if c <= 16 and allow16:
        return 16
    return _ceil(c, 32)


class Block(object):
    def __init__(self, **kw):
        self.kind = 0; self.path = 0
        self.Hi = self.Wi = self.Ci = self.Ho = self.Wo = self.Co = 0
        self.ksize = 3; self.stride = 1; self.pad = 0; self.pool = 0
        self.K = self.N = 0; self.head_kind = 0; self.drop_rate = 0.0
        self.w_off = self.b_off = 0
        # logical description
        self.layer = None            # the weight-carrying layer
        self.ci_logical = self.co_logical = 0
        self.Hs = self.Ws = 0        # stored activation grid (conv)
        self.in_grid = None          # dense after flatten: (Hs, Ws, C_logical, C_pad) of the producing conv block
        self.k_logical = self.n_logical = 0
        self.__dict__.update(kw)


class Plan(object):
    def __init__(self, blocks, nparam, input_shape):
        self.blocks = blocks
        self.nparam = nparam
        self.input_shape = tuple(input_shape)

    # ---- weights: logical list (Keras order: kernel, bias per weighted layer) <-> flat padded buffer ----
    def pack(self, weights):
        flat = np.zeros((self.nparam,), np.float32)
        it = iter(weights)
        for b in self.blocks:
            w = np.asarray(next(it), np.float32); bias = np.asarray(next(it), np.float32)
            if b.kind == KIND_CONV:
                kh, kw, ci, co = w.shape
                wp = np.zeros((kh * kw, b.Ci, b.Co), np.float32)
                wp[:, :ci, :co] = w.reshape(kh * kw, ci, co)
                flat[b.w_off:b.w_off + wp.size] = wp.ravel()
                flat[b.b_off:b.b_off + co] = bias
            else:
                wp = np.zeros((b.K, b.N), np.float32)
                if b.in_grid is not None:
                    hs, ws, c, cp = b.in_grid
                    tmp = np.zeros((hs, ws, cp, b.N), np.float32)
                    tmp[:, :, :c, :b.n_logical] = w.reshape(hs, ws, c, b.n_logical)
                    wp = tmp.reshape(b.K, b.N)
                else:
                    wp[:b.k_logical, :b.n_logical] = w
                flat[b.w_off:b.w_off + wp.size] = wp.ravel()
                flat[b.b_off:b.b_off + b.n_logical] = bias
        return flat

    def unpack(self, flat):
        flat = np.asarray(flat, np.float32)
        out = []
        for b in self.blocks:
            if b.kind == KIND_CONV:
                k2 = b.ksize * b.ksize
                wp = flat[b.w_off:b.w_off + k2 * b.Ci * b.Co].reshape(k2, b.Ci, b.Co)
                out.append(wp[:, :b.ci_logical, :b.co_logical].reshape(b.ksize, b.ksize, b.ci_logical, b.co_logical).copy())
                out.append(flat[b.b_off:b.b_off + b.co_logical].copy())
            else:
                wp = flat[b.w_off:b.w_off + b.K * b.N].reshape(b.K, b.N)
                if b.in_grid is not None:
                    hs, ws, c, cp = b.in_grid
                    w = wp.reshape(hs, ws, cp, b.N)[:, :, :c, :b.n_logical].reshape(hs * ws * c, b.n_logical)
                else:
                    w = wp[:b.k_logical, :b.n_logical]
                out.append(w.copy())
                out.append(flat[b.b_off:b.b_off + b.n_logical].copy())
        return out

    def to_ctypes(self):
        from ._lib import CibBlock
        arr = (CibBlock * len(self.blocks))()
        for i, b in enumerate(self.blocks):
            c = arr[i]
            for f, _ in CibBlock._fields_:
                setattr(c, f, getattr(b, f))
        return arr

    def describe(self):
        rows = []
        for i, b in enumerate(self.blocks):
            if b.kind == KIND_CONV:
                # an un-pooled first layer (1 or 3 input channels) runs on the im2col tensor-core kernels of csrc/conv1_tc.cuh, see
                # conv1_tc_ok() in csrc/runtime.cuh; a pooled one on the tiled CUDA-core kernels (faster there)
                first_tc = (i == 0 and not b.path and b.ksize == 3 and b.stride == 1 and b.Ci in (1, 3) and b.Co in (16, 32, 64)
                            and not b.pool and b.pad in (0, 1))
                rows.append("block %d conv  %dx%dx%d -> %dx%dx%d (stored %dx%d) k%d s%d pad%d pool%d drop%.2f path=%s"
                            % (i, b.Hi, b.Wi, b.Ci, b.Ho, b.Wo, b.Co, b.Hs, b.Ws, b.ksize, b.stride, b.pad, b.pool,
                               b.drop_rate, "tcgen05" if b.path else "tcgen05-im2col" if first_tc else "cuda-core"))
            else:
                rows.append("block %d %s K=%d N=%d drop%.2f path=%s" % (i, "dense" if b.kind == 1 else "head ", b.K, b.N,
                                                                      b.drop_rate, "tcgen05" if b.path else "cuda-core"))
        return "\n".join(rows)


def _tc_conv_ok(ci_mem, co_pad, ksize, stride, first):
    if first or ksize != 3 or stride not in (1, 2):
        return False
    if stride == 2 and ci_mem % 32:          # strided layers run on the per-tap TMA kernels (traversal-stride tensor maps): 32-channel boxes
        return False
    if not (ci_mem == 16 or (ci_mem % 32 == 0 and ci_mem <= 256)):
        return False
    if co_pad % 32 or co_pad > 256:
        return False
    stage = 128 * ci_mem * 4 + ci_mem * co_pad * 4          # forward; dgrad swaps roles (same product)
    stage_d = 128 * co_pad * 4 + ci_mem * co_pad * 4
    if 2 * max(stage, stage_d) + 1024 <= 200 * 1024:
        return True
    # wide layers: the per-tap kernel pipelines 32-channel K chunks instead of whole taps (csrc/conv_tc.cuh, kchunk)
    return ci_mem % 32 == 0 and ci_mem >= 64 and co_pad >= 64


def build_plan(layers, input_shape, force_path=None):
    """``layers``: the model's layer list (InputLayer allowed first).  Returns a :class:`Plan` or raises PlanError."""
    if force_path is None and os.environ.get("CIB_FORCE_PATH") is not None:
        force_path = int(os.environ["CIB_FORCE_PATH"])
    seq = [l for l in layers if not isinstance(l, L.InputLayer)]
    if len(input_shape) != 3:
        raise PlanError("device path expects (H, W, C) inputs")
    H, W, C = input_shape
    if (H * W * C) % 4:
        raise PlanError("sample size must be a multiple of 4 floats")
    blocks = []
    i = 0
    cur_h, cur_w, c_logical, c_mem = H, W, C, C
    # ---------------- conv stack ----------------
    while i < len(seq) and isinstance(seq[i], L.Conv2D):
        conv = seq[i]; i += 1
        if conv.activation != "relu":
            raise PlanError("device path supports relu convolutions")
        if conv.kernel_size[0] != conv.kernel_size[1] or conv.strides[0] != conv.strides[1]:
            raise PlanError("square kernels / strides only")
        k, s = conv.kernel_size[0], conv.strides[0]
        ho, wo, co = conv.compute_output_shape((cur_h, cur_w, c_logical))
        pad = conv.pad_amounts(cur_h, k, s)[0]
        pool, drop = 0, 0.0
        if i < len(seq) and isinstance(seq[i], L.MaxPooling2D):
            mp = seq[i]
            if mp.pool_size != (2, 2) or mp.strides != (2, 2):
                raise PlanError("only 2x2/2 max pooling is fused")
            pool = 1; i += 1
        if i < len(seq) and isinstance(seq[i], L.Dropout):
            drop = seq[i].rate; i += 1
        first = len(blocks) == 0
        co_pad = _ceil(co, 32)
        tc = _tc_conv_ok(c_mem, co_pad, k, s, first)
        if s == 2 and (pool or (blocks and blocks[-1].pool)):
            tc = False          # the strided dgrad (four parity-class launches) has the mask epilogue only, and no fused pooling
        if force_path is not None:
            tc = tc and bool(force_path)
        if not tc:
            # CUDA-core block.  Its consumer may be a tcgen05 block, so keep the channel padding TMA-friendly.
            co_pad = pad_channels(co)
        b = Block(kind=KIND_CONV, path=1 if tc else 0, Hi=cur_h, Wi=cur_w, Ci=c_mem, Ho=ho, Wo=wo, Co=co_pad,
                  ksize=k, stride=s, pad=pad, pool=pool, drop_rate=float(drop), layer=conv,
                  ci_logical=c_logical, co_logical=co)
        b.Hs, b.Ws = (ho // 2, wo // 2) if pool else (ho, wo)
        if b.Hs < 1 or b.Ws < 1:
            raise PlanError("feature map vanished")
        blocks.append(b)
        cur_h, cur_w, c_logical, c_mem = b.Hs, b.Ws, co, co_pad
    if i < len(seq) and isinstance(seq[i], L.Dropout):
        raise PlanError("Dropout placement not supported by the fused device path")
    if i >= len(seq) or not isinstance(seq[i], L.Flatten):
        raise PlanError("expected Flatten after the convolution stack")
    i += 1
    # ---------------- dense stack ----------------
    in_grid = (cur_h, cur_w, c_logical, c_mem) if blocks else None
    k_logical = cur_h * cur_w * c_logical
    k_mem = cur_h * cur_w * c_mem
    dense_layers = []
    while i < len(seq) and isinstance(seq[i], L.Dense):
        d = seq[i]; i += 1
        drop = 0.0
        if i < len(seq) and isinstance(seq[i], L.Dropout):
            drop = seq[i].rate; i += 1
        dense_layers.append((d, drop))
    if i != len(seq):
        raise PlanError("unsupported layer after the dense stack: %s" % type(seq[i]).__name__)
    if len(dense_layers) < 2:
        raise PlanError("device path needs at least one hidden Dense before the output Dense")
    for j, (d, drop) in enumerate(dense_layers):
        last = j == len(dense_layers) - 1
        if last:
            if d.activation not in ("sigmoid", "softmax"):
                raise PlanError("output layer must be sigmoid or softmax")
            if d.units > 16:
                raise PlanError("output layer wider than 16 units")
            b = Block(kind=KIND_HEAD, path=0, K=k_mem, N=d.units, head_kind=0 if d.activation == "sigmoid" else 1,
                      layer=d, k_logical=k_logical, n_logical=d.units, in_grid=None)
        else:
            if d.activation != "relu":
                raise PlanError("hidden Dense layers must be relu")
            n_pad = _ceil(d.units, 32)
            # the GEMM kernels tile N (<= 256 columns per CTA); the loss head reads at most 1024 hidden features
            tc = n_pad <= 1024 and k_mem % 16 == 0
            if force_path is not None:
                tc = tc and bool(force_path)
            b = Block(kind=KIND_HIDDEN, path=1 if tc else 0, K=k_mem, N=n_pad, drop_rate=float(drop), layer=d,
                      k_logical=k_logical, n_logical=d.units, in_grid=in_grid)
            in_grid = None
            k_logical, k_mem = d.units, n_pad
        blocks.append(b)
    # ---------------- flat parameter layout (every tensor starts on a 256-byte boundary) ----------------
    off = 0
    for b in blocks:
        wsize = (b.ksize * b.ksize * b.Ci * b.Co) if b.kind == KIND_CONV else b.K * b.N
        bsize = b.Co if b.kind == KIND_CONV else b.N
        b.w_off = off; off = _ceil(off + wsize, 64)
        b.b_off = off; off = _ceil(off + bsize, 64)
    return Plan(blocks, _ceil(off, 64 * 8), input_shape)
