#!/usr/bin/env python
"""All-reduce bandwidth sweep, 4 KB .. 256 MB: our peer-to-peer kernels (one-shot / two-shot over CUDA-IPC mapped
buffers) against NCCL on the same GPUs.  One process per GPU:

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/gpu_allreduce_sweep.py --out f.json

Times are CUDA events on the launching stream, max over ranks.  busbw = 2(N-1)/N * bytes / t (the NCCL-tests
convention), reported against the NVLink 5 per-direction peak (900 GB/s) -- SURVEY section 8 row 14.
"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from cori_intml_examples_b200.ops import _lib

NVLINK_DIR_GBS = 900.0


def timed(fn, iters, warm):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/allreduce_sweep.json")
    ap.add_argument("--max-mb", type=int, default=256)
    args = ap.parse_args()
    rank, world, lrank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lrank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
    L = _lib.lib()
    _lib.check(L.cib_init(lrank))
    nmax = args.max_mb * (1 << 20) // 4
    h = C.c_void_p()
    _lib.check(L.cib_symbuf_create(C.byref(h), lrank, nmax))
    mine = (C.c_ubyte * 64)()
    _lib.check(L.cib_symbuf_ipc_handle(h, mine))
    gathered = [None] * world
    dist.all_gather_object(gathered, bytes(mine))
    allh = (C.c_ubyte * (64 * world)).from_buffer_copy(b"".join(gathered))
    _lib.check(L.cib_symbuf_attach_ipc(h, rank, world, allh))
    dist.barrier()
    nccl_buf = torch.zeros(nmax, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rows = []
    size = 4096
    while size <= args.max_mb * (1 << 20):
        n = size // 4
        iters = 200 if size <= (1 << 20) else 40 if size <= (32 << 20) else 10
        row = {"bytes": size, "world": world}
        for mode, name in ((1, "one_shot"), (2, "two_shot")):
            if mode == 1 and size > (1 << 20):
                continue
            t = timed(lambda: _lib.check(L.cib_symbuf_allreduce(h, n, 0, mode, C.c_void_p(stream))), iters, 5)
            row[name + "_us"] = t * 1e6
        view = nccl_buf[:n]
        t = timed(lambda: dist.all_reduce(view), iters, 5)
        row["nccl_us"] = t * 1e6
        best = min(row.get("one_shot_us", 1e30), row["two_shot_us"])
        row["ours_us"] = best
        fac = 2.0 * (world - 1) / world
        row["ours_busbw_gbs"] = fac * size / (best * 1e-6) / 1e9
        row["nccl_busbw_gbs"] = fac * size / (row["nccl_us"] * 1e-6) / 1e9
        row["ours_frac_nvlink"] = row["ours_busbw_gbs"] / NVLINK_DIR_GBS
        row["speedup_vs_nccl"] = row["nccl_us"] / best
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        size *= 4
    # correctness spot check against NCCL on the largest size that both ran
    n = 1 << 20
    g = torch.Generator(device="cuda").manual_seed(1234 + rank)
    x = torch.randn(n, device="cuda", generator=g)
    p = C.c_void_p(); _lib.check(L.cib_symbuf_ptr(h, C.byref(p)))

    class _W(object):
        pass
    w = _W(); w.__cuda_array_interface__ = {"shape": (nmax,), "typestr": "<f4", "data": (p.value, False), "version": 3}
    sym = torch.as_tensor(w, device="cuda")
    sym[:n].copy_(x); torch.cuda.synchronize(); dist.barrier()
    _lib.check(L.cib_symbuf_allreduce(h, n, 0, 2, C.c_void_p(stream)))
    ref = x.clone(); dist.all_reduce(ref)
    torch.cuda.synchronize()
    err = float((sym[:n] - ref).abs().max())
    if rank == 0:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump({"world": world, "max_abs_diff_vs_nccl": err, "rows": rows}, f, indent=1)
        print("max |ours - nccl| =", err)
    dist.barrier()
    _lib.check(L.cib_symbuf_destroy(h))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
