#!/usr/bin/env python
"""Run every stand-alone sm_100a kernel against a plain PyTorch fp32 reference, one subprocess per case.

A kernel that traps (bounded mbarrier wait) or faults only kills its own subprocess; the driver loop records the
outcome and moves on.  Results: gpurun_out/kernel_checks.jsonl (one JSON object per case).
Usage:  python scripts/gpu_kernel_checks.py            # all cases
        python scripts/gpu_kernel_checks.py --case gemm_kk_128
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _setup():
    import torch
    torch.cuda.set_device(0)            # PyTorch's notion of the current device and the runtime's (cib_init sets it) must agree
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    from cori_intml_examples_b200.ops import _lib
    L = _lib.lib()
    _lib.check(L.cib_init(0))
    return torch, _lib, L


def _report(name, got, ref, extra=None):
    import torch
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-12
    out = {"case": name, "max_abs_err": err, "ref_max": scale, "rel": err / scale,
           "ok": bool(err / scale < 6e-3 and torch.isfinite(got).all().item())}
    if extra:
        out.update(extra)
    print("RESULT " + json.dumps(out), flush=True)
    return out["ok"]


def case_gemm(name, M, N, K, a_mn, b_mn, splits=1, epi=0):
    torch, _lib, L = _setup()
    g = torch.Generator(device="cuda").manual_seed(1)
    A = torch.randn(M, K, device="cuda", generator=g)
    B = torch.randn(K, N, device="cuda", generator=g)
    ref = A @ B
    Am = A.t().contiguous() if a_mn else A.contiguous()          # MN-major A is stored [K][M]
    Bm = B.contiguous() if b_mn else B.t().contiguous()          # K-major B is stored [N][K]
    Cc = torch.zeros(M, N, device="cuda")
    _lib.check(L.cib_gemm_tf32(_lib.ptr(Am), _lib.ptr(Bm), _lib.ptr(Cc), M, N, K, a_mn, b_mn, splits, epi, None))
    torch.cuda.synchronize()
    return _report(name, Cc, ref)


def case_dense(name, M, N, K):
    torch, _lib, L = _setup()
    g = torch.Generator(device="cuda").manual_seed(5)
    A = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(K, N, device="cuda", generator=g) * 0.05
    b = torch.randn(N, device="cuda", generator=g)
    ref = torch.relu(A @ W + b)
    z = torch.zeros(M, N, device="cuda")
    out = torch.full((M, N), float("nan"), device="cuda")
    ok = True
    for rep in range(2):                 # the scratch must come back zeroed: a second call gives the same answer
        _lib.check(L.cib_dense_relu_fwd(_lib.ptr(A), _lib.ptr(W), _lib.ptr(b), _lib.ptr(out), _lib.ptr(z), M, N, K, None))
        torch.cuda.synchronize()
        ok = _report("%s.call%d" % (name, rep), out, ref) and ok
        ok = ok and float(z.abs().max()) == 0.0
    return ok


def _conv_ref(torch, x, w, bias, pad):
    import torch.nn.functional as F
    # x NHWC, w [3,3,Ci,Co]
    y = F.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), bias, padding=pad)
    return y.permute(0, 2, 3, 1).contiguous()


def case_conv_fwd(name, n, H, W, Ci, Co, same, pool, direct=False):
    torch, _lib, L = _setup()
    import torch.nn.functional as F
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn(n, H, W, Ci, device="cuda", generator=g)
    w = torch.randn(3, 3, Ci, Co, device="cuda", generator=g) * 0.2
    b = torch.randn(Co, device="cuda", generator=g) * 0.1
    pad = 1 if same else 0
    y = torch.relu(_conv_ref(torch, x, w, b, pad))
    if pool:
        y = F.max_pool2d(y.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1).contiguous()
    out = torch.full_like(y, float("nan"))
    if direct:
        _lib.check(L.cib_conv_fwd_direct(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), n, H, W, Ci, Co, 3, 1, pad, int(pool), None))
    else:
        _lib.check(L.cib_conv3x3_fwd_tf32(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), n, H, W, Ci, Co, int(same), int(pool), None))
    torch.cuda.synchronize()
    return _report(name, out, y)


def case_conv_dgrad(name, n, H, W, Ci, Co, same, mode):
    torch, _lib, L = _setup()
    import torch.nn.functional as F
    g = torch.Generator(device="cuda").manual_seed(3)
    pad = 1 if same else 0
    Ho, Wo = (H, W) if same else (H - 2, W - 2)
    dz = torch.randn(n, Ho, Wo, Co, device="cuda", generator=g)
    w = torch.randn(3, 3, Ci, Co, device="cuda", generator=g) * 0.2
    dx = F.conv_transpose2d(dz.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), padding=pad).permute(0, 2, 3, 1).contiguous()
    # stored activation of the previous layer on the (H, W) grid: random zeros + random arg-max bits
    act = torch.rand(n, H, W, Ci, device="cuda", generator=g)
    act = torch.where(act < 0.4, torch.zeros_like(act), act)
    pos = torch.randint(0, 4, act.shape, device="cuda", generator=g, dtype=torch.int32)
    bits = (act.view(torch.int32) & ~3) | pos
    act_s = bits.view(torch.float32)
    scale = 1.25
    gm = torch.where((bits & ~3) != 0, dx * scale, torch.zeros_like(dx))
    if mode == 2:
        ref = gm
        out = torch.full_like(ref, float("nan"))
        Hf, Wf = H, W
    else:
        Hf, Wf = 2 * H, 2 * W
        ref = torch.zeros(n, Hf, Wf, Ci, device="cuda")
        for q in range(4):
            ref[:, (q >> 1)::2, (q & 1)::2, :] = torch.where(pos == q, gm, torch.zeros_like(gm))
        out = torch.full_like(ref, float("nan"))
    _lib.check(L.cib_conv3x3_dgrad_tf32(_lib.ptr(dz), _lib.ptr(w), _lib.ptr(act_s), _lib.ptr(out), n, H, W, Ci, Co, int(same),
                                        mode, Hf, Wf, scale, None))
    torch.cuda.synchronize()
    return _report(name, out, ref)


def case_conv_wgrad(name, n, H, W, Ci, Co, same):
    torch, _lib, L = _setup()
    g = torch.Generator(device="cuda").manual_seed(4)
    pad = 1 if same else 0
    Ho, Wo = (H, W) if same else (H - 2, W - 2)
    x = torch.randn(n, H, W, Ci, device="cuda", generator=g)
    dz = torch.randn(n, Ho, Wo, Co, device="cuda", generator=g)
    dz = torch.where(torch.rand_like(dz) < 0.6, torch.zeros_like(dz), dz)
    w = torch.zeros(3, 3, Ci, Co, device="cuda", requires_grad=True)
    y = _conv_ref(torch, x, w, None, pad)
    (dw_ref,) = torch.autograd.grad(y, w, dz)
    db_ref = dz.sum(dim=(0, 1, 2))
    dW = torch.zeros(3, 3, Ci, Co, device="cuda")
    db = torch.zeros(Co, device="cuda")
    _lib.check(L.cib_conv3x3_wgrad_tf32(_lib.ptr(x), _lib.ptr(dz), _lib.ptr(dW), _lib.ptr(db), n, H, W, Ci, Co, int(same), None))
    torch.cuda.synchronize()
    ok1 = _report(name + ".dW", dW, dw_ref)
    ok2 = _report(name + ".db", db, db_ref)
    return ok1 and ok2


def case_conv1_fwd(name, n, H, W, Ci, Co, same, pool):
    """First-layer tensor-core kernel (im2col producer) against conv2d + relu (+ max_pool2d); with pooling, the arg-max
    bits in the output's two mantissa LSBs must point at a window position holding the maximum."""
    torch, _lib, L = _setup()
    import torch.nn.functional as F
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(n, H, W, Ci, device="cuda", generator=g)
    w = torch.randn(3, 3, Ci, Co, device="cuda", generator=g) * 0.3
    b = torch.randn(Co, device="cuda", generator=g) * 0.1
    pad = 1 if same else 0
    full = torch.relu(_conv_ref(torch, x, w, b, pad))
    y = F.max_pool2d(full.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1).contiguous() if pool else full
    out = torch.full_like(y, float("nan"))
    _lib.check(L.cib_conv1_fwd_tc(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), n, H, W, Ci, Co, pad, int(pool), None))
    torch.cuda.synchronize()
    ok = _report(name, out, y)
    if pool:
        pos = (out.view(torch.int32) & 3).long()
        Hp, Wp = y.shape[1], y.shape[2]
        win = full[:, :2 * Hp, :2 * Wp].reshape(n, Hp, 2, Wp, 2, Co).permute(0, 1, 3, 5, 2, 4).reshape(n, Hp, Wp, Co, 4)
        picked = win.gather(-1, pos.unsqueeze(-1)).squeeze(-1)
        ok = _report(name + ".argmax", picked, y) and ok
    return ok


def case_conv1_wgrad(name, n, H, W, Ci, Co, same, pool):
    torch, _lib, L = _setup()
    import torch.nn.functional as F
    g = torch.Generator(device="cuda").manual_seed(8)
    pad = 1 if same else 0
    Ho, Wo = (H, W) if same else (H - 2, W - 2)
    x = torch.randn(n, H, W, Ci, device="cuda", generator=g)
    if pool:
        Hp, Wp = Ho // 2, Wo // 2
        gp = torch.randn(n, Hp, Wp, Co, device="cuda", generator=g)
        gp = torch.where(torch.rand_like(gp) < 0.4, torch.zeros_like(gp), gp)
        pos = torch.randint(0, 4, (n, Hp, Wp, Co), device="cuda", generator=g)
        act = ((torch.rand(n, Hp, Wp, Co, device="cuda", generator=g) + 0.5).view(torch.int32) & ~3 | pos.int()).view(torch.float32)
        gq = (gp.view(torch.int32) & ~3).view(torch.float32)               # the kernel ignores the two LSBs of g
        dz = torch.zeros(n, Hp, 2, Wp, 2, Co, device="cuda")
        for q in range(4):
            dz[:, :, q >> 1, :, q & 1, :] = torch.where(pos == q, gq, torch.zeros_like(gq))
        dz = dz.reshape(n, 2 * Hp, 2 * Wp, Co)
        if 2 * Hp != Ho or 2 * Wp != Wo:
            dz = F.pad(dz, (0, 0, 0, Wo - 2 * Wp, 0, Ho - 2 * Hp))
        g_in, act_in = gp, act
    else:
        dz = torch.randn(n, Ho, Wo, Co, device="cuda", generator=g)
        dz = torch.where(torch.rand_like(dz) < 0.5, torch.zeros_like(dz), dz)
        g_in, act_in = dz, dz
    w = torch.zeros(3, 3, Ci, Co, device="cuda", requires_grad=True)
    yy = _conv_ref(torch, x, w, None, pad)
    (dw_ref,) = torch.autograd.grad(yy, w, dz)
    db_ref = dz.sum(dim=(0, 1, 2))
    dW = torch.zeros(3, 3, Ci, Co, device="cuda")
    db = torch.zeros(Co, device="cuda")
    _lib.check(L.cib_conv1_wgrad_tc(_lib.ptr(x), _lib.ptr(g_in), _lib.ptr(act_in), _lib.ptr(dW), _lib.ptr(db), n, H, W, Ci, Co, pad, int(pool), None))
    torch.cuda.synchronize()
    ok1 = _report(name + ".dW", dW, dw_ref)
    ok2 = _report(name + ".db", db, db_ref)
    return ok1 and ok2


CASES = {
    "gemm_kk_128": lambda: case_gemm("gemm_kk_128", 128, 128, 128, 0, 0),
    "gemm_kk_k32": lambda: case_gemm("gemm_kk_k32", 128, 64, 32, 0, 0),
    "gemm_kmn": lambda: case_gemm("gemm_kmn", 128, 128, 256, 0, 1),
    "gemm_mnk": lambda: case_gemm("gemm_mnk", 256, 64, 128, 1, 0),
    "gemm_mnmn": lambda: case_gemm("gemm_mnmn", 4096, 128, 128, 1, 1),
    "gemm_splitk": lambda: case_gemm("gemm_splitk", 128, 128, 4096, 0, 1, splits=32, epi=1),
    "gemm_m64": lambda: case_gemm("gemm_m64", 64, 128, 4096, 0, 1, splits=16, epi=1),
    "gemm_n256": lambda: case_gemm("gemm_n256", 128, 256, 512, 0, 0),
    "dense_128_4096": lambda: case_dense("dense_128_4096", 128, 128, 4096),
    "dense_64_2304": lambda: case_dense("dense_64_2304", 64, 32, 2304),
    "conv_fwd_16_32_pool": lambda: case_conv_fwd("conv_fwd_16_32_pool", 4, 32, 32, 16, 32, True, True),
    "conv_fwd_32_64_pool": lambda: case_conv_fwd("conv_fwd_32_64_pool", 4, 16, 16, 32, 64, True, True),
    "conv_fwd_16_32_valid": lambda: case_conv_fwd("conv_fwd_16_32_valid", 3, 26, 26, 16, 32, False, True),
    "conv_fwd_32_32_nopool": lambda: case_conv_fwd("conv_fwd_32_32_nopool", 2, 20, 20, 32, 32, True, False),
    "conv_fwd_64_128": lambda: case_conv_fwd("conv_fwd_64_128", 2, 16, 16, 64, 128, True, True),
    "conv_fwd_128_256_wide": lambda: case_conv_fwd("conv_fwd_128_256_wide", 2, 32, 32, 128, 256, True, False),
    "conv_direct_3_16": lambda: case_conv_fwd("conv_direct_3_16", 4, 64, 64, 3, 16, True, True, direct=True),
    "conv_direct_1_8_valid": lambda: case_conv_fwd("conv_direct_1_8_valid", 4, 28, 28, 1, 8, False, False, direct=True),
    "dgrad_16_32_mask": lambda: case_conv_dgrad("dgrad_16_32_mask", 4, 32, 32, 16, 32, True, 2),
    "dgrad_32_64_unpool": lambda: case_conv_dgrad("dgrad_32_64_unpool", 4, 16, 16, 32, 64, True, 1),
    "dgrad_16_16_valid": lambda: case_conv_dgrad("dgrad_16_16_valid", 3, 26, 26, 16, 16, False, 2),
    "dgrad_128_256_wide": lambda: case_conv_dgrad("dgrad_128_256_wide", 2, 32, 32, 128, 256, True, 2),
    "wgrad_128_256_wide": lambda: case_conv_wgrad("wgrad_128_256_wide", 2, 32, 32, 128, 256, True),
    "wgrad_16_32": lambda: case_conv_wgrad("wgrad_16_32", 8, 32, 32, 16, 32, True),
    "wgrad_32_64": lambda: case_conv_wgrad("wgrad_32_64", 8, 16, 16, 32, 64, True),
    "wgrad_32_32_valid": lambda: case_conv_wgrad("wgrad_32_32_valid", 3, 26, 26, 32, 32, False),
    "wgrad_16_32_valid": lambda: case_conv_wgrad("wgrad_16_32_valid", 3, 26, 26, 16, 32, False),
    "conv1_fwd_3_16_pool": lambda: case_conv1_fwd("conv1_fwd_3_16_pool", 6, 64, 64, 3, 16, True, True),
    "conv1_fwd_1_16_pool": lambda: case_conv1_fwd("conv1_fwd_1_16_pool", 5, 64, 64, 1, 16, True, True),
    "conv1_fwd_1_32_valid": lambda: case_conv1_fwd("conv1_fwd_1_32_valid", 3, 28, 28, 1, 32, False, False),
    "conv1_fwd_3_64_pool": lambda: case_conv1_fwd("conv1_fwd_3_64_pool", 300, 32, 32, 3, 64, True, True),
    "conv1_wgrad_3_16_pool": lambda: case_conv1_wgrad("conv1_wgrad_3_16_pool", 6, 64, 64, 3, 16, True, True),
    "conv1_wgrad_1_16_pool": lambda: case_conv1_wgrad("conv1_wgrad_1_16_pool", 5, 64, 64, 1, 16, True, True),
    "conv1_wgrad_1_32_valid": lambda: case_conv1_wgrad("conv1_wgrad_1_32_valid", 3, 28, 28, 1, 32, False, False),
    "conv1_wgrad_3_64_pool": lambda: case_conv1_wgrad("conv1_wgrad_3_64_pool", 300, 32, 32, 3, 64, True, True),
    "wgrad_64_128": lambda: case_conv_wgrad("wgrad_64_128", 4, 16, 16, 64, 128, True),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case")
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    if args.case:
        ok = CASES[args.case]()
        sys.exit(0 if ok else 3)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    out_path = os.path.join(ROOT, "gpurun_out", "kernel_checks.jsonl")
    names = [n for n in CASES if args.only in n]
    summary = []
    with open(out_path, "w") as f:
        for n in names:
            t0 = time.time()
            try:
                pr = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", n], capture_output=True, text=True, timeout=180)
                rc, so, se = pr.returncode, pr.stdout, pr.stderr
            except subprocess.TimeoutExpired as e:
                rc, so, se = -999, (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or ""), "TIMEOUT"
            results = [json.loads(l[7:]) for l in so.splitlines() if l.startswith("RESULT ")]
            rec = {"case": n, "rc": rc, "secs": round(time.time() - t0, 1), "results": results, "stderr_tail": se[-600:] if rc not in (0, 3) else ""}
            f.write(json.dumps(rec) + "\n"); f.flush()
            summary.append((n, rc, [round(r["rel"], 5) for r in results]))
            print(n, "rc=%d" % rc, [("%.2e" % r["rel"]) for r in results], flush=True)
    bad = [s for s in summary if s[1] != 0]
    print("kernel checks: %d/%d ok" % (len(summary) - len(bad), len(summary)))
    sys.exit(0)


if __name__ == "__main__":
    main()
