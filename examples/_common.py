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
