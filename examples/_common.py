"""Shared plumbing of the example scripts (one script per notebook of the reference).

Every example accepts ``--quick`` (tiny data / few epochs: what the test-suite runs), ``--device auto|cpu|gpu`` and,
where a cluster is involved, ``--engines N`` and ``--cluster-id ID`` (attach to a running cluster instead of starting a
private one).  Data is synthetic unless the reference's files are present (there is no network on the box).
"""
import argparse
import contextlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parser(description, cluster=False, **defaults):
    ap = argparse.ArgumentParser(description=description)
    ap.add_argument("--quick", action="store_true", help="tiny problem sizes (smoke test)")
    ap.add_argument("--device", default="auto", choices=["auto", "cpu", "gpu"])
    ap.add_argument("--out-dir", default=os.environ.get("SCRATCH", "/tmp"))
    ap.add_argument("--record", default=None, help="write wall time, GPU clocks / throttle reasons and the run's results to this JSON file")
    if cluster:
        ap.add_argument("--engines", type=int, default=defaults.get("engines"))
        ap.add_argument("--slots", type=int, default=defaults.get("slots", 1), help="concurrent trials per engine (GPU)")
        ap.add_argument("--cluster-id", default=None, help="attach to this running cluster")
    return ap


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def resolve_device(name):
    if name == "auto":
        return "gpu" if has_gpu() else "cpu"
    return name


@contextlib.contextmanager
def cluster(args, shared_gpus=False, default_engines=2):
    """Yields a connected Client; starts (and stops) a private LocalCluster unless --cluster-id was given."""
    from cori_intml_examples_b200.cluster import Client, LocalCluster
    if args.cluster_id:
        c = Client(cluster_id=args.cluster_id, timeout=60)
        try:
            yield c
        finally:
            c.close()
        return
    use_gpu = resolve_device(args.device) == "gpu"
    n = args.engines or default_engines
    with LocalCluster(n, mode="process", gpus=None if use_gpu else [], slots=args.slots, isolate_gpus=not shared_gpus) as cl:
        c = cl.client()
        try:
            yield c
        finally:
            c.close()


def print_table(rows, columns):
    w = [max(len(str(c)), max((len(_f(r.get(c))) for r in rows), default=0)) for c in columns]
    print("  ".join(str(c).rjust(x) for c, x in zip(columns, w)))
    for r in rows:
        print("  ".join(_f(r.get(c)).rjust(x) for c, x in zip(columns, w)))


def _f(v):
    if isinstance(v, float):
        return "%.4f" % v
    return str(v)


class Recorder(object):
    """``with Recorder(args, name) as rec: ... rec['key'] = value``: wall time of the block, NVML clock samples of every GPU taken
    while it runs (B200_PROFILING.md clocks line: SM MHz median / max, throttle reasons), and whatever the example adds, as one
    JSON file (``--record``).  Without ``--record`` it only measures the wall time."""

    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, args, name):
        import threading
        self.path, self.data = getattr(args, "record", None), {"example": name, "argv": sys.argv[1:]}
        self._stop, self._thread, self._samples = threading.Event(), None, {}
        self._windowed, self._active = False, False

    def __setitem__(self, k, v):
        self.data[k] = v

    def under_load(self):
        """``with rec.under_load():`` -- only clock samples taken inside such blocks count (the training call, not data generation)."""
        rec = self

        class _W(object):
            def __enter__(self_w):
                rec._windowed = True; rec._active = True

            def __exit__(self_w, *a):
                rec._active = False
                return False
        return _W()

    def _poll(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            hs = [nv.nvmlDeviceGetHandleByIndex(i) for i in range(nv.nvmlDeviceGetCount())]
        except Exception:
            return
        for i, h in enumerate(hs):
            self._samples[i] = {"sm": [], "reasons": set(), "max": int(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))}
        while not self._stop.wait(0.02):
            if self._windowed and not self._active:
                continue
            for i, h in enumerate(hs):
                try:
                    self._samples[i]["sm"].append(int(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                    self._samples[i]["reasons"].update(n for n, b in self.REASONS if mask & b)
                except Exception:
                    pass

    def __enter__(self):
        import threading
        import time
        if self.path:
            self._thread = threading.Thread(target=self._poll, daemon=True); self._thread.start()
        self._t0 = time.time()
        return self

    def __exit__(self, et, ev, tb):
        import json
        import time
        self.data["wall_s"] = time.time() - self._t0
        if self._thread is not None:
            self._stop.set(); self._thread.join(2.0)
            self.data["clocks"] = {str(i): {"sm_mhz_median": sorted(v["sm"])[len(v["sm"]) // 2] if v["sm"] else None, "sm_max_mhz": v["max"],
                                            "samples": len(v["sm"]), "reasons": sorted(v["reasons"])} for i, v in self._samples.items()}
        if self.path and self.path != "-" and et is None:
            os.makedirs(os.path.dirname(os.path.abspath(self.path)), exist_ok=True)
            with open(self.path, "w") as f:
                json.dump(self.data, f, indent=1, default=str)
        return False
