#!/usr/bin/env python
"""NVLS (NVSwitch multicast) all-reduce in one process over all visible GPUs: support probe, exactness against a PyTorch sum,
and device-timed comparison with the peer-to-peer kernels (same-process peer access) for 64 KB .. 256 MB.
Writes gpurun_out/nvls_check_<N>gpu.json."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch

import gpu_dist_checks as D
from cori_intml_examples_b200.ops import _lib


def main():
    world = min(torch.cuda.device_count(), int(sys.argv[1]) if len(sys.argv) > 1 else 8)
    L = _lib.lib()
    out = {"world": world, "supported": []}
    for r in range(world):
        _lib.check(L.cib_init(r))
        s = C.c_int(0); _lib.check(L.cib_nvls_supported(r, C.byref(s))); out["supported"].append(s.value)
    nmax = (256 << 20) // 4
    h = C.c_void_p()
    rc = L.cib_nvls_create_local(C.byref(h), world, nmax)
    if rc:
        out["create_error"] = L.cib_last_error().decode()
        print(json.dumps(out)); save(out, world); return
    ptrs = []
    for r in range(world):
        p = C.c_void_p(); _lib.check(L.cib_nvls_ptr(h, r, C.byref(p))); ptrs.append(p.value)
    views = [D.as_tensor(ptrs[r], nmax, r) for r in range(world)]
    # exactness
    checks = []
    for n in (4096, 1 << 20, 5 * (1 << 20) + 4096):
        data = [torch.randn(n, device="cuda:%d" % r) for r in range(world)]
        ref = sum(d.cpu().double() for d in data).float()
        for r in range(world):
            views[r][:n].copy_(data[r])
        for r in range(world):
            torch.cuda.synchronize(r)
        for r in range(world):
            _lib.check(L.cib_nvls_allreduce(h, r, n, 0, None))
        for r in range(world):
            torch.cuda.synchronize(r)
        got = [views[r][:n].cpu() for r in range(world)]
        err = max(float((g - ref).abs().max()) for g in got)
        checks.append({"n": n, "max_abs_err": err, "replicas_identical": all(torch.equal(got[0], g) for g in got[1:]), "ok": err < 1e-4})
    out["checks"] = checks
    # timing against the peer-to-peer kernels (same process, peer access)
    Lp, hs, pp = D.symbufs(world, nmax)
    rows = []
    size = 65536
    while size <= (256 << 20):
        n = size // 4
        iters = 100 if size <= (4 << 20) else 20
        row = {"bytes": size}
        for name, fn in (("nvls", lambda r: L.cib_nvls_allreduce(h, r, n, 0, None)), ("p2p_two_shot", lambda r: Lp.cib_symbuf_allreduce(hs[r], n, 0, 2, None))):
            for _ in range(3):
                for r in range(world):
                    _lib.check(fn(r))
            for r in range(world):
                torch.cuda.synchronize(r)
            ev = []
            for r in range(world):
                with torch.cuda.device(r):
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(torch.cuda.default_stream(r)); ev.append((a, b))
            for _ in range(iters):
                for r in range(world):
                    _lib.check(fn(r))
            for r in range(world):
                with torch.cuda.device(r):
                    ev[r][1].record(torch.cuda.default_stream(r))
            for r in range(world):
                torch.cuda.synchronize(r)
            row[name + "_us"] = max(a.elapsed_time(b) for a, b in ev) * 1e3 / iters
        fac = 2.0 * (world - 1) / world
        row["nvls_busbw_gbs"] = fac * size / (row["nvls_us"] * 1e-6) / 1e9
        row["p2p_busbw_gbs"] = fac * size / (row["p2p_two_shot_us"] * 1e-6) / 1e9
        rows.append(row)
        print(json.dumps(row), flush=True)
        size *= 8
    out["rows"] = rows
    _lib.check(L.cib_nvls_destroy(h))
    print(json.dumps({k: out[k] for k in ("world", "supported", "checks")}))
    save(out, world)


def save(out, world):
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "nvls_check_%dgpu.json" % world), "w"), indent=1)


if __name__ == "__main__":
    main()
