"""CPU oracle backend: plain PyTorch fp32 ops + autograd, NumPy optimizers with the Keras update rules.

This is (a) the numerics reference every CUDA kernel is tested against, (b) the backend used on machines
without a GPU (driver CPU tests, gloo multi-process tests), and (c) the timed CPU baseline for
BASELINE.json config 1 (MNIST CNN single-process training on CPU).
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import layers as L

EPS = 1e-7


def keras_loss_and_acc(probs, y, loss_name):
    """Per-sample Keras losses on *probabilities* (clip to [eps, 1-eps]) and per-sample accuracy."""
    if loss_name == "categorical_crossentropy":
        p = probs / probs.sum(dim=-1, keepdim=True)
        p = torch.clamp(p, EPS, 1.0 - EPS)
        loss = -(y * torch.log(p)).sum(dim=-1)
        acc = (probs.argmax(dim=-1) == y.argmax(dim=-1)).float()
    elif loss_name == "binary_crossentropy":
        p = torch.clamp(probs, EPS, 1.0 - EPS)
        loss = -(y * torch.log(p) + (1.0 - y) * torch.log(1.0 - p)).mean(dim=-1)
        acc = (torch.round(probs) == y).float().mean(dim=-1)
    else:
        raise ValueError("unsupported loss %r" % (loss_name,))
    return loss, acc


class CpuBackend(object):
    name = "cpu"

    def __init__(self, model):
        self.model = model
        self.layers = model.layers
        self.weights = [np.array(w, np.float32) for w in model._initial_weights()]
        self.slots = model.optimizer.init_slots(self.weights) if model.optimizer is not None else []
        self.rng = torch.Generator().manual_seed(model.seed)

    # ------------------------------------------------------------------ weights
    def get_weights(self):
        return [w.copy() for w in self.weights]

    def set_weights(self, weights):
        for dst, src in zip(self.weights, weights):
            dst[...] = np.asarray(src, np.float32).reshape(dst.shape)

    def get_slots(self):
        return [[a.copy() for a in s] for s in self.slots]

    def set_slots(self, slots):
        for dst, src in zip(self.slots, slots):
            for d, s in zip(dst, src):
                d[...] = np.asarray(s, np.float32).reshape(d.shape)

    def reset_optimizer(self):
        self.slots = self.model.optimizer.init_slots(self.weights)

    def sync_optimizer_to_host(self):
        pass

    def set_lr(self, lr):
        pass

    # ------------------------------------------------------------------ math
    def forward(self, x, params, training):
        """x: torch NHWC float32.  Returns output probabilities."""
        h = x
        nchw = False
        it = iter(params)
        for layer in self.layers:
            if isinstance(layer, L.InputLayer):
                continue
            if isinstance(layer, L.Conv2D):
                if not nchw:
                    h = h.permute(0, 3, 1, 2); nchw = True
                w = next(it); b = next(it)
                kh, kw = layer.kernel_size
                (pt, pb), (pl, pr) = layer.pad_amounts(h.shape[2], kh, layer.strides[0]), layer.pad_amounts(h.shape[3], kw, layer.strides[1])
                if pt or pb or pl or pr:
                    h = F.pad(h, (pl, pr, pt, pb))
                h = F.conv2d(h, w.permute(3, 2, 0, 1), b, stride=layer.strides)
                h = self._act(h, layer.activation, ch_dim=1)
            elif isinstance(layer, L.MaxPooling2D):
                h = F.max_pool2d(h, layer.pool_size, layer.strides)
            elif isinstance(layer, L.Dropout):
                if training and layer.rate > 0:
                    keep = (torch.rand(h.shape, generator=self.rng) >= layer.rate).to(h.dtype)
                    h = h * keep / (1.0 - layer.rate)
            elif isinstance(layer, L.Flatten):
                if nchw:
                    h = h.permute(0, 2, 3, 1); nchw = False
                h = h.reshape(h.shape[0], -1)
            elif isinstance(layer, L.Dense):
                w = next(it); b = next(it)
                h = self._act(h @ w + b, layer.activation, ch_dim=-1)
            else:
                raise TypeError("unsupported layer %r" % (layer,))
        if nchw:
            h = h.permute(0, 2, 3, 1)
        return h

    @staticmethod
    def _act(h, name, ch_dim):
        if name == "relu":
            return torch.relu(h)
        if name == "sigmoid":
            return torch.sigmoid(h)
        if name == "softmax":
            return torch.softmax(h, dim=ch_dim)
        if name in (None, "linear"):
            return h
        raise ValueError("unsupported activation %r" % (name,))

    def _tensors(self, requires_grad):
        ts = [torch.from_numpy(w) for w in self.weights]
        if requires_grad:
            ts = [t.clone().requires_grad_(True) for t in ts]
        return ts

    def loss_and_grads(self, x, y, sample_weight=None):
        """Batch-mean loss, accuracy and gradients (list of arrays in weight order) -- no update."""
        params = self._tensors(True)
        probs = self.forward(torch.from_numpy(np.ascontiguousarray(x, np.float32)), params, True)
        yt = torch.from_numpy(np.ascontiguousarray(y, np.float32)).reshape(probs.shape)
        loss, acc = keras_loss_and_acc(probs, yt, self.model.loss)
        if sample_weight is not None:
            sw = torch.from_numpy(np.asarray(sample_weight, np.float32))
            mean_loss = (loss * sw).sum() / sw.sum()
            mean_acc = (acc * sw).sum() / sw.sum()
        else:
            mean_loss, mean_acc = loss.mean(), acc.mean()
        grads = torch.autograd.grad(mean_loss, params)
        return float(mean_loss.detach()), float(mean_acc.detach()), [g.numpy() for g in grads]

    def train_on_batch(self, x, y):
        loss, acc, grads = self.loss_and_grads(x, y)
        grads = self.model._reduce_gradients(grads)
        self.model.optimizer.apply(self.weights, grads, self.slots)
        return loss, acc

    def fit_stream(self, batches):
        return [self.train_on_batch(np.asarray(x), np.asarray(y))[0] for x, y in batches]

    def test_on_batch(self, x, y):
        with torch.no_grad():
            probs = self.forward(torch.from_numpy(np.ascontiguousarray(x, np.float32)), self._tensors(False), False)
            yt = torch.from_numpy(np.ascontiguousarray(y, np.float32)).reshape(probs.shape)
            loss, acc = keras_loss_and_acc(probs, yt, self.model.loss)
        return float(loss.sum()), float(acc.sum())

    def predict_on_batch(self, x):
        with torch.no_grad():
            return self.forward(torch.from_numpy(np.ascontiguousarray(x, np.float32)), self._tensors(False), False).numpy()

    # ------------------------------------------------------------------ epoch-level API (shared with the GPU backend)
    def fit_epoch(self, x, y, batch_size, perm, on_batch=None):
        n = len(perm)
        tot_loss = tot_acc = 0.0
        seen = 0
        for bi, start in enumerate(range(0, n, batch_size)):
            idx = perm[start:start + batch_size]
            loss, acc = self.train_on_batch(x[idx], y[idx])
            tot_loss += loss * len(idx); tot_acc += acc * len(idx); seen += len(idx)
            if on_batch is not None:
                on_batch(bi, len(idx), tot_loss / seen, tot_acc / seen)
        return tot_loss / seen, tot_acc / seen

    def evaluate(self, x, y, batch_size):
        n = len(x)
        tl = ta = 0.0
        for start in range(0, n, batch_size):
            l, a = self.test_on_batch(x[start:start + batch_size], y[start:start + batch_size])
            tl += l; ta += a
        return tl / n, ta / n

    def predict(self, x, batch_size):
        outs = [self.predict_on_batch(x[s:s + batch_size]) for s in range(0, len(x), batch_size)]
        return np.concatenate(outs, axis=0)

    def close(self):
        pass
