#!/usr/bin/env python
"""HPO throughput (trials/hour) on the local GPUs -- BASELINE.json configs 3 and 4:

  * ``--workload mnist``: random search of mnist.build_model, 64 trials load-balanced over the GPUs, each trial =
    16 epochs x 49,800 training samples + 10,200 validation samples, batch 128 (DistHPO_mnist.ipynb:136-145,187-190)
  * ``--workload rpv --widgets``: 32 RPV trials (16 epochs x 64,000 / 32,000, batch 64, DistHPO_rpv.ipynb:90-106) monitored
    live by a headless ParamSpanWidget fed by IPyParallelLogger.

trials/hour = completed trials / wall time from first submit to last result.  ``--slots k`` packs k
concurrent trials per GPU (each its own engine object, stream and CUDA graphs inside one process / CUDA context).
"""
import argparse
import functools
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def mnist_trial(n_train, n_test, n_epochs, batch_size, device, h1, h2, h3, dropout, optimizer, checkpoint_file=None):
    import time as _t
    t0 = _t.time()
    from cori_intml_examples_b200.models import mnist
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    from cori_intml_examples_b200.nn.callbacks import ModelCheckpoint
    x_train, y_train, _, _ = mnist.load_data(n_train, n_test)
    t1 = _t.time()
    model = mnist.build_model(h1=h1, h2=h2, h3=h3, dropout=dropout, optimizer=optimizer, device=device)
    # every reference trial writes model_<i>.h5 after each epoch (DistHPO_mnist.ipynb:184-185,248)
    cbs = [IPyParallelLogger()] + ([ModelCheckpoint(checkpoint_file)] if checkpoint_file else [])
    h = model.fit(x_train, y_train, batch_size=batch_size, epochs=n_epochs, validation_split=0.17, callbacks=cbs, verbose=0)
    out = dict(h.history)
    out["_t_data"], out["_t_fit"] = t1 - t0, _t.time() - t1
    return out


def rpv_trial(n_train, n_valid, n_epochs, batch_size, device, conv_sizes, fc_sizes, lr, dropout, optimizer, checkpoint_file=None):
    import time as _t
    t0 = _t.time()
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    (xt, yt, _), (xv, yv, _), _ = rpv.load_dataset("/nonexistent", n_train=n_train, n_valid=n_valid, n_test=0)
    t1 = _t.time()
    model = rpv.build_model(xt.shape[1:], conv_sizes=conv_sizes, fc_sizes=fc_sizes, dropout=dropout, optimizer=optimizer, lr=lr, device=device)
    h = rpv.train_model(model, xt, yt, xv, yv, batch_size=batch_size, n_epochs=n_epochs, verbose=0, callbacks=[IPyParallelLogger()],
                        checkpoint_file=checkpoint_file)       # per-epoch model_<i>.h5 as in DistHPO_rpv.ipynb
    out = dict(h.history)
    out["_t_data"], out["_t_fit"] = t1 - t0, _t.time() - t1
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="mnist", choices=["mnist", "rpv"])
    ap.add_argument("--trials", type=int, default=None)
    ap.add_argument("--epochs", type=int, default=16)
    ap.add_argument("--engines", type=int, default=None, help="default: one per GPU")
    ap.add_argument("--slots", type=int, default=1)
    ap.add_argument("--device", default="gpu")
    ap.add_argument("--widgets", action="store_true")
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the reference's sample counts")
    ap.add_argument("--out", default=None)
    ap.add_argument("--channels", type=int, default=3, help="RPV image channels (BASELINE.json: 64x64x3)")
    ap.add_argument("--label-noise", type=float, default=0.05, help="fraction of synthetic labels flipped: validation accuracy stays below 1")
    ap.add_argument("--no-checkpoints", action="store_true", help="skip the per-epoch model_<i>.h5 writes every reference trial does")
    ap.add_argument("--ckpt-dir", default=os.path.join(os.environ.get("SCRATCH", "/tmp"), "cib_hpo_ckpt_%d" % os.getpid()))
    a = ap.parse_args()
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import _common
    from cori_intml_examples_b200 import hpo
    from cori_intml_examples_b200.cluster import LocalCluster
    from cori_intml_examples_b200.monitor import ParamSpanWidget, ModelPlot
    if a.workload == "mnist":
        n = a.trials or 64
        trials = hpo.sample_mnist_space(n, seed=0)
        func = functools.partial(mnist_trial, int(60000 * a.scale), int(10000 * a.scale), a.epochs, 128, a.device)
        per_trial = "%d epochs x %d train + %d val samples, batch 128" % (a.epochs, int(int(60000 * a.scale) * 0.83), int(60000 * a.scale) - int(int(60000 * a.scale) * 0.83))
    else:
        n = a.trials or 32
        trials = hpo.sample_rpv_space(n, seed=0)
        func = functools.partial(rpv_trial, int(64000 * a.scale), int(32000 * a.scale), a.epochs, 64, a.device)
        per_trial = "%d epochs x %d train + %d val samples, batch 64" % (a.epochs, int(64000 * a.scale), int(32000 * a.scale))
    gpus = None if a.device == "gpu" else []
    if not a.no_checkpoints:
        os.makedirs(a.ckpt_dir, exist_ok=True)
        for i, t in enumerate(trials):
            t["checkpoint_file"] = os.path.join(a.ckpt_dir, "model_%d.h5" % i)
    a.record = None
    eng_env = {"RPV_CHANNELS": str(a.channels), "CIB_SYNTH_LABEL_NOISE": str(a.label_noise)}
    with LocalCluster(a.engines, mode="process", gpus=gpus, slots=a.slots, env=eng_env) as cl, _common.Recorder(argparse.Namespace(record="-"), "hpo_bench") as rec:
        c = cl.client()
        n_eng = len(c.ids)
        # warm the engines (imports, CUDA context, library load) outside the timed region, as a long-lived cluster would be
        c[:].execute("import cori_intml_examples_b200.models.mnist, cori_intml_examples_b200.models.rpv\n"
                     "import torch\n_ = torch.zeros(1).cuda() if torch.cuda.is_available() else None", silent=True)
        t0 = time.time()
        polls = 0
        if a.widgets:
            keys = list(trials[0].keys())
            params = {k: [t[k] for t in trials] for k in keys}
            vis = functools.partial(ModelPlot, y=['loss', 'val_loss', 'acc', 'val_acc'], x='epoch', xlim=(0, a.epochs - 1), ylim=(0, 1))
            psw = ParamSpanWidget(func, vis, params, client=c, interval=1.0)
            psw.submit_computations()
            while not psw.all_done():
                time.sleep(0.2); polls += 1
            wall = time.time() - t0
            psw.poll_once()
            frame_txt = psw.render_text()
            statuses = psw.get_models_status()
            val = [r.get("val_acc") for r in psw.rows]
            psw.close()
            done = sum(s == "Ended Training" for s in statuses)
            extra = {"widget_rows_ended": done, "table_head": frame_txt.splitlines()[:6]}
            times = []
        else:
            rs = hpo.RandomSearch(c.load_balanced_view(), func, trials).submit()
            rs.wait()
            wall = time.time() - t0
            hist = rs.histories()
            done = rs.n_done()
            times = rs.task_times()
            val = [max(h["val_acc"]) for h in hist]
            extra = {"t_data_mean": float(np.mean([h["_t_data"] for h in hist])), "t_fit_mean": float(np.mean([h["_t_fit"] for h in hist])),
                     "task_time_mean": float(np.nanmean(times)), "task_time_max": float(np.nanmax(times))}
        res = {"workload": a.workload, "trials": n, "completed": int(done), "engines": n_eng, "slots": a.slots, "wall_s": wall,
               "trials_per_hour": 3600.0 * done / wall, "per_trial": per_trial, "device": a.device, "widgets": bool(a.widgets),
               "best_val_acc": float(np.nanmax([v for v in val if v is not None])) if any(v is not None for v in val) else None}
        res.update(extra)
        rec._stop.set(); rec._thread.join(2.0)
        res["clocks"] = {str(i): {"sm_mhz_median": sorted(v["sm"])[len(v["sm"]) // 2] if v["sm"] else None, "sm_max_mhz": v["max"], "samples": len(v["sm"]),
                                  "reasons": sorted(v["reasons"])} for i, v in rec._samples.items()}
        res["checkpoints"] = (not a.no_checkpoints); res["label_noise"] = a.label_noise
        if a.workload == "rpv":
            res["input"] = "64x64x%d" % a.channels
        if not a.no_checkpoints:
            files = [f for f in os.listdir(a.ckpt_dir) if f.endswith(".h5")]
            res["checkpoint_files"] = len(files); res["checkpoint_bytes"] = sum(os.path.getsize(os.path.join(a.ckpt_dir, f)) for f in files)
        res["val_acc_all"] = [None if v is None else round(float(v), 4) for v in val]
        print(json.dumps(res))
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            with open(a.out, "w") as f:
                json.dump(res, f, indent=1)
        c.close()


if __name__ == "__main__":
    main()
