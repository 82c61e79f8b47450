#!/usr/bin/env python
"""Print in-kernel timelines (globaltimer stamps) for selected kernels at the headline shapes."""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from cori_intml_examples_b200.ops import _lib
L = _lib.lib(); _lib.check(L.cib_init(0)); P = _lib.ptr
B = 128; r = lambda *s: torch.randn(*s, device="cuda")
def trace(name, fn):
    for _ in range(2): _lib.check(fn())
    torch.cuda.synchronize()
    _lib.check(L.cib_ktrace_enable(1))
    _lib.check(fn()); torch.cuda.synchronize()
    buf = (C.c_ulonglong * 64)(); _lib.check(L.cib_ktrace_read(buf)); _lib.check(L.cib_ktrace_enable(0))
    for cta in range(2):
        v = list(buf[cta * 32:(cta + 1) * 32]); t0 = v[0]
        print("%-28s cta%d " % (name, cta) + " ".join("%d:%.1f" % (i, (x - t0) / 1e3) for i, x in enumerate(v) if x), flush=True)
A = r(B, 4096); W = r(4096, 128); Z = torch.zeros(B, 128, device="cuda"); DZ = r(B, 128); DX = torch.empty(B, 4096, device="cuda")
trace("dense1_fwd", lambda: L.cib_gemm_tf32(P(A), P(W), P(Z), B, 128, 4096, 0, 1, 32, 1, None))
trace("dense1_dgrad", lambda: L.cib_gemm_tf32(P(DZ), P(W), P(DX), B, 4096, 128, 0, 0, 1, 0, None))
x1 = r(B, 32, 32, 16); w1 = r(3, 3, 16, 32) * .1; b32 = r(32); o1 = torch.empty(B, 16, 16, 32, device="cuda")
trace("conv2_fwd", lambda: L.cib_conv3x3_fwd_tf32(P(x1), P(w1), P(b32), P(o1), B, 32, 32, 16, 32, 1, 1, None))
x2 = r(B, 16, 16, 32); w2 = r(3, 3, 32, 64) * .1; b64 = r(64); o2 = torch.empty(B, 8, 8, 64, device="cuda")
trace("conv3_fwd", lambda: L.cib_conv3x3_fwd_tf32(P(x2), P(w2), P(b64), P(o2), B, 16, 16, 32, 64, 1, 1, None))
