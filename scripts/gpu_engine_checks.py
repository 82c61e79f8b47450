#!/usr/bin/env python
"""End-to-end checks of the sm_100a engine against the CPU oracle (gradients, optimizer steps, short trainings).
One subprocess per case; results in gpurun_out/engine_checks.jsonl.
Models: rpv.build_model (rpv.py:38-72), mnist.build_model (mnist.py:44-59), legacy strided model (Train_rpv.ipynb:204-218)."""
import argparse, json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def rel(a, b):
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-12))


def cos(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30))


def build(kind, device, opt="SGD", dropout=0.0, lr=0.05):
    from cori_intml_examples_b200.models import mnist, rpv
    # the initialisers draw from NumPy's global stream: fixed seed per (kind, device) so that a case is the same in every run
    # (unseeded, a threshold was missed once in a few dozen runs)
    np.random.seed((sum(ord(c) for c in kind) * 31 + (0 if device == "cpu" else 7)) % (2 ** 31))
    from cori_intml_examples_b200.nn import optimizers
    o = optimizers.get(opt, lr=lr) if opt != "default" else "Adam"
    if kind == "rpv3":
        m = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=dropout, optimizer=opt, lr=lr, device=device)
    elif kind == "rpv1":
        m = rpv.build_model((64, 64, 1), [8, 16, 32], [64], dropout=dropout, optimizer=opt, lr=lr, device=device)
    elif kind == "rpv_odd":
        m = rpv.build_model((64, 64, 1), [13, 34, 70], [153], dropout=dropout, optimizer=opt, lr=lr, device=device)
    elif kind == "mnist":
        m = mnist.build_model(8, 16, 32, dropout, optimizer=o, device=device)
    elif kind == "mnist_big":
        m = mnist.build_model(64, 64, 128, dropout, optimizer=o, device=device)
    elif kind == "legacy":
        m = rpv.build_legacy_model((32, 32, 1), 8, 16, 16, 16, 32, optimizer=o, device=device)
    elif kind == "legacy_wide":            # every strided layer on the tcgen05 path (32+ input channels), a K-chunked wide layer in between
        m = rpv.build_legacy_model((32, 32, 1), 32, 64, 128, 128, 320, optimizer=o, device=device)
    return m


def data_for(kind, n, seed=0):
    from cori_intml_examples_b200.utils import synthetic, np_utils
    if kind.startswith("mnist"):
        x, y = synthetic.mnist_like(n, seed)
        return x, np_utils.to_categorical(y, 10)
    ch = 3 if kind == "rpv3" else 1
    size = 32 if kind.startswith("legacy") else 64
    x, y, _ = synthetic.rpv_like(n, seed, size=size, channels=ch)
    return x, y.reshape(-1, 1)


def case_grads(kind, B=32):
    """Gradients of one batch: GPU engine (tcgen05 path) vs CPU oracle."""
    import ctypes as C
    from cori_intml_examples_b200.ops import _lib
    x, y = data_for(kind, B)
    cpu = build(kind, "cpu"); gpu = build(kind, "cuda")
    gpu.seed = cpu.seed
    w = cpu.get_weights(); gpu.set_weights(w)
    loss_c, acc_c, g_c = cpu.backend.loss_and_grads(x, y)
    be = gpu.backend
    be._ensure(B)
    be._bind(x, y, None)
    _lib.check(be.lib.cib_engine_eval_steps(be.engine, 1, None))          # gather the batch into the engine buffers
    _lib.check(be.lib.cib_engine_sync(be.engine))
    be._upload(1, np.zeros(be.plan.nparam, np.float32))
    _lib.check(be.lib.cib_engine_reset_metrics(be.engine))
    _lib.check(be.lib.cib_engine_forward_backward(be.engine, 1, 1))
    m = be.metrics()
    g_g = be.plan.unpack(be._download(1))
    out = {"case": "grads_" + kind, "loss_cpu": loss_c, "loss_gpu": m.step_loss, "acc_cpu": acc_c, "acc_gpu": m.step_acc,
           "plan": be.plan.describe().split("\n")}
    rels = [rel(a, b) for a, b in zip(g_g, g_c)]
    out["grad_rel"] = rels
    # tf32 rounding flips a few ReLU / arg-max decisions on sparse inputs; perturbing the fp32 oracle's weights by
    # 5e-4 moves its own conv gradients by 2-11 % (measured), so direction (cosine) is the meaningful criterion here.
    # The kernels themselves are checked tightly in isolation by gpu_kernel_checks.py.
    out["grad_cos"] = [cos(a, b) for a, b in zip(g_g, g_c)]
    out["ok"] = bool(min(out["grad_cos"]) > 0.99 and max(rels) < 0.25 and abs(loss_c - m.step_loss) < 2e-3 * max(1, abs(loss_c)))
    return out


def case_step(kind, opt, B=32, steps=3):
    """`steps` optimizer steps on the same batch: weights GPU (graph replay) vs CPU oracle (no dropout)."""
    x, y = data_for(kind, B)
    cpu = build(kind, "cpu", opt, lr={"SGD": 0.05, "Adam": 0.001, "Nadam": 0.001, "Adadelta": 1.0}[opt])
    gpu = build(kind, "cuda", opt, lr={"SGD": 0.05, "Adam": 0.001, "Nadam": 0.001, "Adadelta": 1.0}[opt])
    w_init = cpu.get_weights()
    gpu.set_weights(w_init)
    for _ in range(steps):
        cpu.backend.train_on_batch(x, y)
    perm = np.arange(B, dtype=np.int32)
    for _ in range(steps):
        gpu.backend.fit_epoch(x, y, B, perm)
    wc, wg = cpu.get_weights(), gpu.get_weights()
    rels = [float(np.percentile(np.abs(a - b), 99.9) / (np.abs(a).max() + 1e-12)) for a, b in zip(wc, wg)]
    dcos = [cos(a - c, b - c) for a, b, c in zip(wc, wg, w_init)]
    gpu.backend.sync_optimizer_to_host()
    out = {"case": "step_%s_%s" % (kind, opt), "w_rel": rels, "delta_cos": dcos, "iters": gpu.optimizer.iterations,
           "launches": gpu.backend.launch_counts()}
    # biases start at zero and take sign-like Adam/Adadelta steps: their direction is noisier under tf32 (see case_grads)
    out["ok"] = bool(min(dcos[0::2]) > 0.97 and min(dcos[1::2]) > 0.85 and gpu.optimizer.iterations == steps)
    return out


def case_train(kind, opt="Adam", n=2048, B=64, epochs=3, dropout=0.2):
    x, y = data_for(kind, n); xv, yv = data_for(kind, 512, seed=5)
    gpu = build(kind, "cuda", opt, dropout=dropout, lr=0.001 if opt == "Adam" else 1.0)
    t0 = time.time()
    h = gpu.fit(x, y, batch_size=B, epochs=epochs, validation_data=(xv, yv), verbose=0)
    dt = time.time() - t0
    p = gpu.predict(xv[:100]); ev = gpu.evaluate(xv, yv, verbose=0)
    out = {"case": "train_%s_%s" % (kind, opt), "history": h.history, "secs": dt, "pred_shape": list(p.shape), "eval": ev,
           "finite": bool(np.isfinite(p).all())}
    out["ok"] = bool(h.history["loss"][-1] < h.history["loss"][0] and np.isfinite(h.history["val_loss"]).all()
                     and abs(ev[0] - h.history["val_loss"][-1]) < 1e-3)
    return out


CASES = {}
for k in ("rpv3", "rpv1", "mnist", "legacy", "legacy_wide", "rpv_odd", "mnist_big"):
    CASES["grads_" + k] = (case_grads, (k,))
for o in ("SGD", "Adam", "Nadam", "Adadelta"):
    CASES["step_rpv3_" + o] = (case_step, ("rpv3", o))
CASES["step_mnist_Adam"] = (case_step, ("mnist", "Adam"))
CASES["train_rpv3"] = (case_train, ("rpv3",))
CASES["train_mnist"] = (case_train, ("mnist", "Adadelta"))
CASES["train_legacy"] = (case_train, ("legacy", "Adam", 512, 32, 2, 0.0))


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--case"); ap.add_argument("--only", default="")
    a = ap.parse_args()
    if a.case:
        fn, args = CASES[a.case]
        out = fn(*args)
        print("RESULT " + json.dumps(out), flush=True)
        sys.exit(0 if out["ok"] else 3)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", "engine_checks.jsonl")
    nbad = 0
    with open(path, "w") as f:
        for n in [c for c in CASES if a.only in c]:
            t0 = time.time()
            try:
                pr = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", n], capture_output=True, text=True, timeout=300)
                rc, so, se = pr.returncode, pr.stdout, pr.stderr
            except subprocess.TimeoutExpired:
                rc, so, se = -999, "", "TIMEOUT"
            res = [json.loads(l[7:]) for l in so.splitlines() if l.startswith("RESULT ")]
            f.write(json.dumps({"case": n, "rc": rc, "secs": round(time.time() - t0, 1), "results": res, "stderr_tail": se[-1500:] if rc != 0 else ""}) + "\n"); f.flush()
            brief = ""
            if res:
                r = res[0]
                brief = str({k: (np.round(v, 5).tolist() if isinstance(v, list) and v and isinstance(v[0], float) else v) for k, v in r.items() if k in ("grad_cos", "delta_cos", "loss_cpu", "loss_gpu", "iters", "launches", "secs", "eval")})
            print(n, "rc=%d" % rc, brief, flush=True)
            if rc != 0:
                nbad += 1
                print("   stderr:", se[-600:].replace("\n", " | "), flush=True)
    print("engine checks: %d bad" % nbad)


if __name__ == "__main__":
    main()
