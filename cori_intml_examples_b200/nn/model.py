"""Sequential / functional Model with Keras 2.2 training semantics on top of the two backends.

fit(): per-epoch reshuffle, last partial batch kept, running sample-weighted means of loss/acc, ``validation_split``
takes the *last* fraction before shuffling (60000 * 0.17 -> 49800 / 10200, DistHPO_mnist.ipynb:293), validation in
inference mode at each epoch end, history keys ``loss / acc / val_loss / val_acc`` (Keras 2.2 naming, hard-coded
in hpo_widgets.py:178 and mlextras.py:15-18).

Reference call sites: compile mnist.py:57-58 / rpv.py:69-71; fit rpv.py:103-106, DistHPO_mnist.ipynb:187-190;
evaluate train_rpv.py:83-85; predict DistTrain_rpv.ipynb:407-408; summary train_rpv.py:66; save/load_model
DistHPO_mnist.ipynb:185,542.
"""
import sys
import time

import numpy as np

from . import callbacks as cbks
from . import layers as L
from . import optimizers as optimizers_mod

_model_counter = {"sequential": 0, "model": 0}


def _auto_model_name(prefix):
    _model_counter[prefix] += 1
    return "%s_%d" % (prefix, _model_counter[prefix])


def _loss_name(loss):
    if callable(loss):
        loss = getattr(loss, "__name__", str(loss))
    if loss not in ("categorical_crossentropy", "binary_crossentropy"):
        raise ValueError("unsupported loss %r (categorical_crossentropy / binary_crossentropy)" % (loss,))
    return loss


class Model(object):
    """Functional model over a linear chain: ``Model(inputs=Input(...), outputs=layer(...(inputs)))``."""

    def __init__(self, inputs=None, outputs=None, name=None, _layers=None):
        self.name = name or _auto_model_name("model" if _layers is None else "sequential")
        if _layers is not None:
            self.layers = list(_layers)
        else:
            chain = outputs.chain()
            if chain[0] is not inputs.layer:
                raise ValueError("outputs are not connected to inputs")
            self.layers = chain
        self.optimizer = None
        self.loss = None
        self.metrics_names = ["loss"]
        self.backend = None
        self.stop_training = False
        self.seed = int(np.random.randint(0, 2 ** 31 - 1))
        self.history = None
        self._dist = None            # set by parallel.dist.DistributedOptimizer
        self._built = False
        self._device_pref = "auto"
        if self.layers:
            self._build()

    # ------------------------------------------------------------------ construction
    def _build(self):
        first = self.layers[0]
        if isinstance(first, L.InputLayer):
            shape = first.shape
        elif first.input_shape_arg is not None:
            shape = first.input_shape_arg
        else:
            raise ValueError("the first layer needs input_shape=...")
        self.input_shape = tuple(int(s) for s in shape)
        cur = self.input_shape
        for layer in self.layers:
            layer.build(cur)
            cur = layer.output_shape
        self.output_shape = cur
        self._built = True

    def _initial_weights(self):
        rng = np.random.RandomState(self.seed)
        ws = []
        for layer in self.layers:
            ws.extend(layer.init_weights(rng))
        return ws

    @property
    def weighted_layers(self):
        return [l for l in self.layers if l.has_weights]

    def count_params(self):
        return int(sum(l.count_params() for l in self.layers))

    # ------------------------------------------------------------------ compile / backend
    def compile(self, optimizer="Adam", loss="categorical_crossentropy", metrics=None, device="auto"):
        """``optimizer`` is a name (library defaults, mnist.py:57), an instance (rpv.py:62) or a
        ``DistributedOptimizer`` wrapper (rpv.py:65).  ``device``: 'auto' | 'cpu' | 'cuda' | 'cuda:N'."""
        from ..parallel import dist as dist_mod
        if isinstance(optimizer, dist_mod.DistributedOptimizer):
            self._dist = optimizer
            optimizer = optimizer.optimizer
        self.optimizer = optimizers_mod.get(optimizer)
        self.loss = _loss_name(loss)
        self.metrics_names = ["loss"] + (["acc"] if metrics else [])
        self._device_pref = device
        old = self.backend
        weights = old.get_weights() if old is not None else None
        self.backend = self._make_backend(device)
        if weights is not None:
            self.backend.set_weights(weights)
            old.close()
        if self._dist is not None:
            self._dist.bind(self)
        return self

    def _make_backend(self, device):
        import torch
        device = str(device).lower()
        if device == "gpu" or device.startswith("gpu:"):
            device = "cuda" + device[3:]
        if not (device in ("auto", "cpu", "cuda") or (device.startswith("cuda:") and device[5:].isdigit())):
            raise ValueError("unknown device %r (use 'auto', 'cpu', 'gpu'/'cuda' or 'cuda:N')" % (device,))
        if device.startswith("cuda") and not torch.cuda.is_available():
            raise RuntimeError("device %r requested but no CUDA GPU is visible" % (device,))
        want_gpu = device.startswith("cuda") or (device == "auto" and torch.cuda.is_available())
        if want_gpu:
            from ..ops.plan import PlanError
            from .backend_gpu import GpuBackend
            idx = int(device.split(":")[1]) if ":" in device else None
            try:
                return GpuBackend(self, idx)
            except PlanError as e:
                if device != "auto":
                    raise
                sys.stderr.write("[cori_intml_examples_b200] layer chain not supported by the fused device path (%s); "
                                 "using the CPU oracle backend\n" % e)
        from .backend_cpu import CpuBackend
        return CpuBackend(self)

    def _require_compiled(self):
        if self.backend is None:
            # inference-only use (e.g. a freshly loaded model without training config)
            self.backend = self._make_backend(self._device_pref)

    def _reduce_gradients(self, grads):
        """CPU-oracle data parallelism: average gradients across ranks (gloo / shared-memory collective)."""
        if self._dist is None:
            return grads
        return self._dist.allreduce_gradients(grads)

    # ------------------------------------------------------------------ weights
    def get_weights(self):
        self._require_compiled()
        return self.backend.get_weights()

    def set_weights(self, weights):
        self._require_compiled()
        ws = list(weights)
        shapes = [s for l in self.layers for s in l.weight_shapes()]
        if len(ws) != len(shapes):
            raise ValueError("expected %d weight arrays, got %d" % (len(shapes), len(ws)))
        for w, s in zip(ws, shapes):
            if tuple(np.shape(w)) != tuple(s):
                raise ValueError("weight shape %s does not match %s" % (np.shape(w), s))
        self.backend.set_weights(ws)

    def get_lr(self):
        return self.optimizer.lr

    def set_lr(self, lr):
        self.optimizer.lr = float(lr)
        self.backend.set_lr(lr)

    # ------------------------------------------------------------------ training
    def fit(self, x=None, y=None, batch_size=32, epochs=1, verbose=1, callbacks=None, validation_split=0.0,
            validation_data=None, shuffle=True, initial_epoch=0):
        if self.optimizer is None:
            raise RuntimeError("compile() the model before fit()")
        x = np.asarray(x) if not hasattr(x, "shape") else x
        y = np.asarray(y) if not hasattr(y, "shape") else y
        y = y.reshape(len(y), -1)
        val_x = val_y = None
        if validation_data is not None:
            val_x, val_y = validation_data[0], np.asarray(validation_data[1]).reshape(len(validation_data[1]), -1)
        elif validation_split and 0.0 < validation_split < 1.0:
            split_at = int(len(x) * (1.0 - validation_split))
            x, val_x = x[:split_at], x[split_at:]
            y, val_y = y[:split_at], y[split_at:]
        n = len(x)
        do_val = val_x is not None
        history = cbks.History()
        cb_list = list(callbacks or []) + [history]      # Keras 2.2 order: History last, so it records 'lr' and averaged metrics
        params = {"batch_size": batch_size, "epochs": epochs, "steps": None, "samples": n, "verbose": verbose,
                  "do_validation": do_val, "metrics": ["loss", "acc"] + (["val_loss", "val_acc"] if do_val else [])}
        cl = cbks.CallbackList(cb_list, self, params)
        self.history = history
        self.stop_training = False
        self.backend.push_hyper() if hasattr(self.backend, "push_hyper") else None
        if verbose and self._is_chief():
            if do_val:
                print("Train on %d samples, validate on %d samples" % (n, len(val_x)))
            else:
                print("Train on %d samples" % n)
        cl.call("on_train_begin", {})
        rng = np.random.RandomState(self.seed + 7919)
        for epoch in range(initial_epoch, epochs):
            if self.stop_training:
                break
            cl.call("on_epoch_begin", epoch, {})
            if verbose and self._is_chief():
                print("Epoch %d/%d" % (epoch + 1, epochs))
            t0 = time.time()
            perm = rng.permutation(n) if shuffle else np.arange(n)
            on_batch = _Progbar(n, batch_size).update if (verbose == 1 and self._is_chief()) else None
            loss, acc = self.backend.fit_epoch(x, y, batch_size, perm.astype(np.int32), on_batch)
            logs = {"loss": float(loss), "acc": float(acc)}
            if do_val:
                vl, va = self.backend.evaluate(val_x, val_y, batch_size)
                logs["val_loss"], logs["val_acc"] = float(vl), float(va)
            self.backend.sync_optimizer_to_host()
            cl.call("on_epoch_end", epoch, logs)
            if verbose and self._is_chief():
                dt = time.time() - t0
                body = " - ".join("%s: %.4f" % (k, logs[k]) for k in ("loss", "acc", "val_loss", "val_acc") if k in logs)
                if verbose == 1:
                    print("\r%d/%d [==============================] - %ds %dus/step - %s" % (n, n, dt, int(dt * 1e6 / max(n, 1)), body))
                else:
                    print(" - %ds - %s" % (dt, body))
        cl.call("on_train_end", {})
        return history

    def _is_chief(self):
        return True

    def train_on_batch(self, x, y):
        """One optimizer step on a host batch; returns [loss, acc]."""
        y = np.asarray(y) if not hasattr(y, "shape") else y
        loss, acc = self.backend.train_on_batch(x, y.reshape(len(y), -1))
        return [loss, acc]

    def fit_stream(self, batches):
        """Train on an iterable of host (x, y) batches (e.g. pinned tensors from a data loader); returns the per-step
        losses.  On the GPU backend the host->device copies overlap the previous step (see GpuBackend.fit_stream)."""
        if self.optimizer is None:
            raise RuntimeError("compile() the model before fit_stream()")
        return self.backend.fit_stream(batches)

    def evaluate(self, x=None, y=None, batch_size=32, verbose=1):
        self._require_compiled()
        if self.loss is None:
            raise RuntimeError("evaluate() needs a compiled model (loss)")
        y = np.asarray(y).reshape(len(y), -1)
        loss, acc = self.backend.evaluate(x, y, batch_size)
        if verbose == 1:
            print("%d/%d [==============================] - 0s" % (len(x), len(x)))
        return [float(loss), float(acc)] if len(self.metrics_names) > 1 else float(loss)

    def predict(self, x, batch_size=32, verbose=0):
        self._require_compiled()
        return self.backend.predict(x, batch_size)

    def predict_classes(self, x, batch_size=32, verbose=0):
        p = self.predict(x, batch_size)
        if p.shape[-1] > 1:
            return p.argmax(axis=-1)
        return (p > 0.5).astype("int32")

    def predict_proba(self, x, batch_size=32, verbose=0):
        return self.predict(x, batch_size)

    # ------------------------------------------------------------------ summary / config / persistence
    def summary(self, line_length=65, print_fn=None):
        print_fn = print_fn or print
        pos = [int(line_length * p) for p in (0.45, 0.85, 1.0)]

        def row(fields):
            line = ""
            for f, p in zip(fields, pos):
                line += str(f)
                line = line[:p] + " " * (p - len(line))
            print_fn(line)
        print_fn("_" * line_length)
        row(["Layer (type)", "Output Shape", "Param #"])
        print_fn("=" * line_length)
        for i, layer in enumerate(self.layers):
            row(["%s (%s)" % (layer.name, layer.class_name), str((None,) + tuple(layer.output_shape)), layer.count_params()])
            print_fn(("=" if i == len(self.layers) - 1 else "_") * line_length)
        total = self.count_params()
        print_fn("Total params: {:,}".format(total))
        print_fn("Trainable params: {:,}".format(total))
        print_fn("Non-trainable params: 0")
        print_fn("_" * line_length)

    def get_config(self):
        layer_cfgs = [{"class_name": l.class_name, "config": l.get_config()} for l in self.layers]
        if isinstance(self, Sequential):
            return {"name": self.name, "layers": [c for c in layer_cfgs if c["class_name"] != "InputLayer"]}
        # functional: linear chain with inbound nodes
        out = []
        prev = None
        for l, c in zip(self.layers, layer_cfgs):
            c = dict(c, name=l.name, inbound_nodes=[] if prev is None else [[[prev.name, 0, 0, {}]]])
            out.append(c); prev = l
        return {"name": self.name, "layers": out, "input_layers": [[self.layers[0].name, 0, 0]],
                "output_layers": [[self.layers[-1].name, 0, 0]]}

    def save(self, filepath, overwrite=True, include_optimizer=True):
        from . import saving
        saving.save_model(self, filepath, include_optimizer=include_optimizer)

    def save_weights(self, filepath):
        from . import saving
        saving.save_weights(self, filepath)

    def load_weights(self, filepath):
        from . import saving
        saving.load_weights(self, filepath)


class Sequential(Model):
    """``Sequential()`` + ``.add(layer)`` (reference mnist.py:47-56)."""

    def __init__(self, layers=None, name=None):
        self._pending = []
        Model.__init__(self, name=name or _auto_model_name("sequential"), _layers=[])
        for l in layers or []:
            self.add(l)

    def add(self, layer):
        self.layers.append(layer)
        if self.layers[0].input_shape_arg is not None or isinstance(self.layers[0], L.InputLayer):
            self._build()


class _Progbar(object):
    """Minimal Keras-like progress line for verbose=1."""

    def __init__(self, total, batch_size):
        self.total, self.batch_size, self.seen, self.last = total, batch_size, 0, 0.0

    def update(self, batch_index, batch_len, loss, acc):
        self.seen = min(self.total, self.seen + batch_len)
        now = time.time()
        if now - self.last < 0.2 and self.seen < self.total:
            return
        self.last = now
        width = 30
        filled = int(width * self.seen / max(self.total, 1))
        bar = "=" * max(filled - 1, 0) + (">" if filled < width else "=") + "." * (width - filled)
        sys.stdout.write("\r%d/%d [%s] - loss: %.4f - acc: %.4f" % (self.seen, self.total, bar, loss, acc))
        sys.stdout.flush()
