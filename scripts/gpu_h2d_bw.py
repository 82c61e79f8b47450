#!/usr/bin/env python
"""Pinned host -> device copy bandwidth of this box (the ceiling of bench.py's end-to-end number, which moves 6.29 MB per step):
one copy per step on one stream, or the same bytes split into chunks issued on several streams."""
import json, torch
res = {}
total = int(6.29e6) // 4
h = torch.empty(total, dtype=torch.float32).pin_memory(); d = torch.empty(total, device="cuda")
for nstreams, nchunks in ((1, 1), (1, 4), (2, 2), (2, 4), (4, 4), (4, 8), (8, 8)):
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    cs = (total + nchunks - 1) // nchunks
    def step():
        for c in range(nchunks):
            with torch.cuda.stream(streams[c % nstreams]):
                d[c * cs:(c + 1) * cs].copy_(h[c * cs:(c + 1) * cs], non_blocking=True)
    for _ in range(3): step()
    torch.cuda.synchronize()
    import time
    t0 = time.perf_counter()
    for _ in range(50): step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / 50
    res["%d streams x %d chunks" % (nstreams, nchunks)] = {"ms_per_6.29MB": round(ms, 4), "GB_per_s": round(total * 4 / ms / 1e6, 2)}
for mb in (64, 256):
    n = int(mb * 1e6) // 4
    hh = torch.empty(n, dtype=torch.float32).pin_memory(); dd = torch.empty(n, device="cuda")
    dd.copy_(hh, non_blocking=True); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): dd.copy_(hh, non_blocking=True)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    res["one copy of %d MB" % mb] = {"ms": round(ms, 4), "GB_per_s": round(n * 4 / ms / 1e6, 2)}
print(json.dumps({"h2d_pinned": res}, indent=1))
