#!/usr/bin/env python
"""Device-time every stand-alone kernel at the headline RPV shapes (per-GPU batch 128), CUDA events, warm L2.
Writes gpurun_out/kernel_bench.json.  --only <substr> restricts the cases (used under ncu)."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cori_intml_examples_b200.ops import _lib

ap = argparse.ArgumentParser(); ap.add_argument("--only", default=""); ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
L = _lib.lib(); _lib.check(L.cib_init(0))
B = 128
dev = "cuda"
r = lambda *s: torch.randn(*s, device=dev)
P = _lib.ptr


def bits(act):
    return ((act.view(torch.int32) & ~3) | torch.randint(0, 4, act.shape, device=dev, dtype=torch.int32)).view(torch.float32)


cases = {}
x0 = r(B, 64, 64, 3); w0 = r(3, 3, 3, 16) * .1; b16 = r(16); o0 = torch.empty(B, 32, 32, 16, device=dev)
cases["conv1_direct_fwd 3->16 64x64"] = (lambda: L.cib_conv_fwd_direct(P(x0), P(w0), P(b16), P(o0), B, 64, 64, 3, 16, 3, 1, 1, 1, None), 2 * B * 64 * 64 * 27 * 16)
x1 = r(B, 32, 32, 16); w1 = r(3, 3, 16, 32) * .1; b32 = r(32); o1 = torch.empty(B, 16, 16, 32, device=dev)
cases["conv2_tc_fwd 16->32 32x32"] = (lambda: L.cib_conv3x3_fwd_tf32(P(x1), P(w1), P(b32), P(o1), B, 32, 32, 16, 32, 1, 1, None), 2 * B * 32 * 32 * 144 * 32)
x2 = r(B, 16, 16, 32); w2 = r(3, 3, 32, 64) * .1; b64 = r(64); o2 = torch.empty(B, 8, 8, 64, device=dev)
cases["conv3_tc_fwd 32->64 16x16"] = (lambda: L.cib_conv3x3_fwd_tf32(P(x2), P(w2), P(b64), P(o2), B, 16, 16, 32, 64, 1, 1, None), 2 * B * 16 * 16 * 288 * 64)
dz3 = r(B, 16, 16, 64); a2 = bits(torch.rand(B, 16, 16, 32, device=dev)); dz2o = torch.empty(B, 32, 32, 32, device=dev)
cases["conv3_tc_dgrad ->unpool 32x32x32"] = (lambda: L.cib_conv3x3_dgrad_tf32(P(dz3), P(w2), P(a2), P(dz2o), B, 16, 16, 32, 64, 1, 1, 32, 32, 1.0, None), 2 * B * 16 * 16 * 288 * 64)
dz2 = r(B, 32, 32, 32); a1 = bits(torch.rand(B, 32, 32, 16, device=dev)); g1 = torch.empty(B, 32, 32, 16, device=dev)
cases["conv2_tc_dgrad ->mask 32x32x16"] = (lambda: L.cib_conv3x3_dgrad_tf32(P(dz2), P(w1), P(a1), P(g1), B, 32, 32, 16, 32, 1, 2, 32, 32, 1.0, None), 2 * B * 32 * 32 * 144 * 32)
dW2 = torch.zeros(3, 3, 32, 64, device=dev); db2 = torch.zeros(64, device=dev)
cases["conv3_tc_wgrad 32->64"] = (lambda: L.cib_conv3x3_wgrad_tf32(P(x2), P(dz3), P(dW2), P(db2), B, 16, 16, 32, 64, 1, None), 2 * B * 16 * 16 * 288 * 64)
dW1 = torch.zeros(3, 3, 16, 32, device=dev); db1 = torch.zeros(32, device=dev)
x1s = torch.cat([x1.reshape(-1), torch.zeros(64, device=dev)])[: x1.numel()].reshape(x1.shape)
cases["conv2_tc_wgrad 16->32"] = (lambda: L.cib_conv3x3_wgrad_tf32(P(x1), P(dz2), P(dW1), P(db1), B, 32, 32, 16, 32, 1, None), 2 * B * 32 * 32 * 144 * 32)
A = r(B, 4096); W = r(4096, 128); Z = torch.zeros(B, 128, device=dev)
cases["dense1_fwd splitK 128x128x4096"] = (lambda: L.cib_gemm_tf32(P(A), P(W), P(Z), B, 128, 4096, 0, 1, 32, 1, None), 2 * B * 128 * 4096)
DZ = r(B, 128); DWd = torch.empty(4096, 128, device=dev)
cases["dense1_wgrad 4096x128x128"] = (lambda: L.cib_gemm_tf32(P(A), P(DZ), P(DWd), 4096, 128, B, 1, 1, 1, 0, None), 2 * B * 128 * 4096)
DX = torch.empty(B, 4096, device=dev)
cases["dense1_dgrad 128x4096x128"] = (lambda: L.cib_gemm_tf32(P(DZ), P(W), P(DX), B, 4096, 128, 0, 0, 1, 0, None), 2 * B * 128 * 4096)

out = {}
for name, (fn, flops) in cases.items():
    if a.only not in name:
        continue
    for _ in range(3):
        _lib.check(fn())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(a.iters):
        _lib.check(fn())
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / a.iters
    out[name] = {"us": us, "tflops": flops / us / 1e6}
    print("%-40s %8.1f us  %7.1f TFLOP/s" % (name, us, flops / us / 1e6), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "kernel_bench%s.json" % ("_" + a.only.split()[0] if a.only else "")), "w"), indent=1)
