"""GPU backend: drives the sm_100a engine (libcib200.so) -- device-resident data, CUDA-graph replayed steps.

Reference behaviour reproduced here: ``model.fit`` per-epoch semantics (rpv.py:103-106, DistHPO_mnist.ipynb:187-190: shuffled
batches, last partial batch kept, sample-weighted running means), ``evaluate`` / ``predict`` in inference mode
(train_rpv.py:83-85, DistTrain_rpv.ipynb:407-408); SURVEY 2.4.1.

One engine = one model replica on one GPU with a fixed batch size.  ``fit_epoch`` uploads the epoch's
permutation (256 KB for 64k samples), replays the captured train graph ``ceil(n / B)`` times without any host
synchronisation and reads the device-side metric accumulators once.  Nothing on this path calls a PyTorch op,
cuDNN, cuBLAS or NCCL; PyTorch is used as the tensor/allocator interop for the dataset buffers only.
A missing library is a hard error here (never a silent fallback to the CPU oracle).
"""
import ctypes as C
import os

import numpy as np
import torch

from ..ops import _lib
from ..ops.plan import build_plan


# Device-resident copies of host datasets, shared by every model of the process (an HPO engine trains many models on the same arrays):
# key -> (x_dev, y_dev, n, (x, y), bytes).  Keyed on the host buffers' identity AND a cheap content fingerprint, so an array that was
# modified in place between two fits is uploaded again instead of silently training on the stale copy; least-recently-used entries are
# dropped once the cache holds more than CIB_DATASET_CACHE_GB (default 16) per process.
import collections
import threading

_DEVICE_DATASETS = collections.OrderedDict()
_DEVICE_DATASETS_LOCK = threading.Lock()


def _fingerprint(a):
    """A few thousand strided samples of the buffer: cheap, and catches in-place edits (shuffles, re-normalisation, new labels)."""
    if a is None:
        return 0
    if isinstance(a, torch.Tensor):
        flat = a.reshape(-1)
        step = max(1, flat.numel() // 4096)
        return (int(a.data_ptr()), float(flat[::step][:4096].double().sum()))
    flat = np.asarray(a).reshape(-1)
    step = max(1, flat.size // 4096)
    return (int(flat.__array_interface__["data"][0]), float(np.asarray(flat[::step][:4096], np.float64).sum()))


class GpuBackend(object):
    name = "b200"

    def __init__(self, model, device=None):
        self.model = model
        self.lib = _lib.lib()
        if device is None:
            # engines that share GPU visibility (data-parallel jobs across engines) are told their GPU through CIB_DEVICE
            env = os.environ.get("CIB_DEVICE")
            device = int(env) if env not in (None, "") else torch.cuda.current_device()
        self.device = int(device)
        _lib.check(self.lib.cib_init(self.device))
        self.plan = build_plan(model.layers, model.input_shape)
        self.engine = None
        self.batch = None
        self._pending_weights = model._initial_weights()
        self._pending_slots = None
        self._datasets = {}          # keys of the process-wide device dataset cache this backend is (or was recently) bound to
        self._comm = None            # (comm handle, rank, mode)
        self.use_graph = True

    # ------------------------------------------------------------------ engine life cycle
    def _ensure(self, batch):
        if self.engine is not None and self.batch == batch:
            return
        if self.engine is not None:
            self._gather_slots()            # collective under data parallelism: every rank changes its batch size at the same point
            self._pending_weights = self.get_weights()
            self._pending_slots = self.get_slots()
            self._destroy()
        h, w, c = self.model.input_shape
        eng = C.c_void_p()
        opt = self.model.optimizer
        kind = opt.kind if opt is not None else 1
        torch.cuda.set_device(self.device)
        _lib.check(self.lib.cib_engine_create(C.byref(eng), self.plan.to_ctypes(), len(self.plan.blocks), int(batch),
                                              h, w, c, self.plan.nparam, kind, 1 if self.use_graph else 0))
        self.engine, self.batch = eng, int(batch)
        self._upload(0, self.plan.pack(self._pending_weights))
        if self._pending_slots is not None:
            for i, s in enumerate(self._pending_slots[:2]):
                self._upload(2 + i, self.plan.pack(s))
        self._pending_weights = None; self._pending_slots = None
        self.push_hyper()
        if self._comm is not None:
            _lib.check(self.lib.cib_engine_attach_comm(self.engine, *self._comm))
        elif getattr(self.model, "_dist", None) is not None:
            self.model._dist.setup_engine(self)       # collective: every rank creates its engine at the same point

    def _destroy(self):
        if self.engine is not None:
            self.lib.cib_engine_destroy(self.engine)
            self.engine = None
            self._dist_comm_attached = False

    def close(self):
        self._destroy()
        self._datasets.clear()

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    def _gather_slots(self):
        """Data parallelism only: the two-shot fused update leaves each rank with its own slice of the optimizer moments; pull the
        other slices from their owners (NVLink) so that checkpoints / re-created engines / broadcasts see complete slots.  Collective."""
        if self.engine is not None and getattr(self, "_dist_comm_attached", False):
            _lib.check(self.lib.cib_engine_gather_slots(self.engine))

    def _dist_barrier(self):
        """Host barrier before a rank's first collective of an epoch: dataset upload and graph capture take different times on
        different ranks, and the device-side waits of the collectives are bounded."""
        if getattr(self, "_dist_comm_attached", False):
            from ..parallel import dist
            dist.barrier()

    def _upload(self, which, flat):
        flat = np.ascontiguousarray(flat, np.float32)
        _lib.check(self.lib.cib_engine_upload(self.engine, which, flat.ctypes.data_as(C.c_void_p), flat.size))

    def _download(self, which):
        flat = np.empty((self.plan.nparam,), np.float32)
        _lib.check(self.lib.cib_engine_download(self.engine, which, flat.ctypes.data_as(C.c_void_p), flat.size))
        return flat

    def push_hyper(self):
        opt = self.model.optimizer
        if opt is None or self.engine is None:
            return
        h = opt.hyper()
        hy = _lib.CibHyper(h["lr"], h["beta1"], h["beta2"], h["eps"], h["rho"], h["momentum"], h["schedule_decay"],
                           h["decay"], h["m_schedule"], int(h["iterations"]), int(self.model.seed) & 0xFFFFFFFF)
        _lib.check(self.lib.cib_engine_set_hyper(self.engine, C.byref(hy)))

    def metrics(self):
        m = _lib.CibMetrics()
        _lib.check(self.lib.cib_engine_get_metrics(self.engine, C.byref(m)))
        return m

    def sync_optimizer_to_host(self):
        if self.engine is None or self.model.optimizer is None:
            return
        m = self.metrics()
        self.model.optimizer.iterations = int(m.iterations)
        if hasattr(self.model.optimizer, "m_schedule"):
            self.model.optimizer.m_schedule = float(m.m_schedule)

    def set_lr(self, lr):
        if self.engine is not None:
            _lib.check(self.lib.cib_engine_set_lr(self.engine, float(lr)))

    # ------------------------------------------------------------------ weights / optimizer slots
    def get_weights(self):
        if self.engine is None:
            return [np.array(w, np.float32) for w in self._pending_weights]
        return self.plan.unpack(self._download(0))

    def set_weights(self, weights):
        if self.engine is None:
            self._pending_weights = [np.array(w, np.float32) for w in weights]
        else:
            self._upload(0, self.plan.pack(weights))

    def get_slots(self):
        n = self.model.optimizer.n_slots if self.model.optimizer is not None else 0
        if self.engine is None:
            if self._pending_slots is not None:
                return self._pending_slots
            return [[np.zeros_like(w) for w in self.get_weights()] for _ in range(n)]
        return [self.plan.unpack(self._download(2 + i)) for i in range(n)]

    def set_slots(self, slots):
        if self.engine is None:
            self._pending_slots = [[np.array(a, np.float32) for a in s] for s in slots]
        else:
            for i, s in enumerate(slots[:2]):
                self._upload(2 + i, self.plan.pack(s))

    def reset_optimizer(self):
        if self.engine is not None:
            z = np.zeros((self.plan.nparam,), np.float32)
            self._upload(2, z); self._upload(3, z)
        self._pending_slots = None
        self.push_hyper()

    # ------------------------------------------------------------------ data
    def _dataset(self, x, y):
        """Device-resident copy of (x, y), cached per host array so every epoch re-uses it."""
        # keyed on the buffers, not on the Python objects: fit() makes fresh reshape / validation-split views of the same memory each call
        key = (self.device, tuple(x.shape), None if y is None else tuple(np.shape(y)), self.plan.blocks[-1].N, _fingerprint(x), _fingerprint(y))
        with _DEVICE_DATASETS_LOCK:
            hit = _DEVICE_DATASETS.get(key)
            if hit is not None:
                _DEVICE_DATASETS.move_to_end(key)
                return hit[:4]
        dev = torch.device("cuda", self.device)
        with _lib.exclusive():      # allocation + synchronisation: keep them away from other trials' graph captures
            xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, np.float32))
            xd = xt.to(dev, torch.float32).contiguous()
            n = xd.shape[0]
            if y is None:
                yd = torch.zeros((n, self.plan.blocks[-1].N), device=dev)
            else:
                yt = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y, np.float32))
                yd = yt.to(dev, torch.float32).reshape(n, -1).contiguous()
            if yd.shape[1] != self.plan.blocks[-1].N:
                raise ValueError("label width %d does not match the model output %d" % (yd.shape[1], self.plan.blocks[-1].N))
            torch.cuda.current_stream(dev).synchronize()
        nbytes = xd.numel() * 4 + yd.numel() * 4
        entry = (xd, yd, n, (x, y), nbytes)
        cap = float(os.environ.get("CIB_DATASET_CACHE_GB", "16")) * 2 ** 30
        with _DEVICE_DATASETS_LOCK:
            _DEVICE_DATASETS[key] = entry
            total = sum(e[4] for e in _DEVICE_DATASETS.values())
            while total > cap and len(_DEVICE_DATASETS) > 1:
                _, old = _DEVICE_DATASETS.popitem(last=False)
                total -= old[4]
        self._datasets[key] = entry              # keeps the tensors alive while this engine may still be bound to them
        if len(self._datasets) > 4:
            self._datasets.pop(next(iter(self._datasets)))
        return entry[:4]

    def _bind(self, x, y, perm=None):
        xd, yd, n, _ = self._dataset(x, y)
        _lib.check(self.lib.cib_engine_set_dataset(self.engine, _lib.ptr(xd), _lib.ptr(yd), n, yd.shape[1]))
        if perm is not None:
            p = np.ascontiguousarray(perm, np.int32)
            if len(p) != n:
                # training on a subset (validation_split): build a full-length table whose head is the subset
                raise ValueError("permutation must cover the bound dataset")
            _lib.check(self.lib.cib_engine_set_perm(self.engine, p.ctypes.data_as(C.c_void_p), n))
        return n

    # ------------------------------------------------------------------ epoch-level API
    def fit_epoch(self, x, y, batch_size, perm, on_batch=None):
        self._ensure(batch_size)
        n = self._bind(x, y, perm)
        _lib.check(self.lib.cib_engine_reset_metrics(self.engine))
        steps = -(-n // self.batch)
        self._dist_barrier()
        if on_batch is None:
            _lib.check(self.lib.cib_engine_train_steps(self.engine, steps))
        else:
            chunk = max(1, steps // 20)
            done = 0
            while done < steps:
                k = min(chunk, steps - done)
                _lib.check(self.lib.cib_engine_train_steps(self.engine, k))
                done += k
                m = self.metrics()
                s = max(m.sample_sum, 1e-20)
                on_batch(done - 1, k * self.batch, m.loss_sum / s, m.correct_sum / s)
        self._gather_slots()
        m = self.metrics()
        s = max(m.sample_sum, 1e-20)
        return m.loss_sum / s, m.correct_sum / s

    def evaluate(self, x, y, batch_size):
        self._ensure(self.batch or batch_size)
        n = self._bind(x, y, None)
        _lib.check(self.lib.cib_engine_reset_metrics(self.engine))
        _lib.check(self.lib.cib_engine_eval_steps(self.engine, -(-n // self.batch), None))
        m = self.metrics()
        s = max(m.sample_sum, 1e-20)
        return m.loss_sum / s, m.correct_sum / s

    def predict(self, x, batch_size):
        self._ensure(self.batch or batch_size)
        n = self._bind(x, None, None)
        steps = -(-n // self.batch)
        width = self.plan.blocks[-1].N
        out = torch.empty((steps * self.batch, width), device=torch.device("cuda", self.device))
        _lib.check(self.lib.cib_engine_eval_steps(self.engine, steps, _lib.ptr(out)))
        _lib.check(self.lib.cib_engine_sync(self.engine))
        return out[:n].cpu().numpy()

    def train_on_batch(self, x, y):
        """Host-fed step (pinned-memory H2D + step + D2H of the loss); used by bench.py's end-to-end number."""
        self._ensure(len(x))
        xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, np.float32))
        yt = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y, np.float32))
        if not hasattr(self, "_loss_pin"):
            self._loss_pin = torch.zeros(2, dtype=torch.float32).pin_memory()
        _lib.check(self.lib.cib_engine_train_batch_host(self.engine, _lib.ptr(xt), _lib.ptr(yt), len(x), _lib.ptr(self._loss_pin)))
        _lib.check(self.lib.cib_engine_sync(self.engine))
        return float(self._loss_pin[0]), float(self._loss_pin[1])

    def fit_stream(self, batches):
        """Host-fed training over an iterable of (x, y) batches, pipelined two deep: while step i computes, the H2D copy of step
        i+1 is in flight.  Every step's {loss, accuracy} is written by the loss head straight into a pinned host ring (zero-copy
        device-to-host, no copy node between two step graphs); the ring is drained at every synchronisation.  Returns the losses."""
        losses, pending, it0, ring, entries = [], 0, None, None, 0
        for x, y in batches:
            self._ensure(len(x))
            if ring is None:
                rp, ne = C.POINTER(C.c_float)(), C.c_int(0)
                _lib.check(self.lib.cib_engine_loss_ring(self.engine, C.byref(rp), C.byref(ne)))
                entries = ne.value
                ring = np.ctypeslib.as_array(rp, shape=(entries, 2))
                it0 = int(self.metrics().iterations)          # synchronises once, before the first step
            xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, np.float32))
            yt = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y, np.float32))
            _lib.check(self.lib.cib_engine_train_batch_host(self.engine, _lib.ptr(xt), _lib.ptr(yt), len(xt), None))
            pending += 1
            if pending == 512:
                _lib.check(self.lib.cib_engine_sync(self.engine))
                losses += [float(ring[(it0 + k) % entries, 0]) for k in range(pending)]
                it0 += pending; pending = 0
        if ring is not None:
            self._gather_slots()
            _lib.check(self.lib.cib_engine_sync(self.engine))
            losses += [float(ring[(it0 + k) % entries, 0]) for k in range(pending)]
        return losses

    def test_on_batch(self, x, y):
        l, a = self.evaluate(x, y, len(x))
        return l * len(x), a * len(x)

    def attach_comm(self, comm, rank, mode=0):
        self._comm = (comm, int(rank), int(mode))
        if self.engine is not None:
            _lib.check(self.lib.cib_engine_attach_comm(self.engine, comm, int(rank), int(mode)))

    def launch_counts(self):
        t, v = C.c_int(0), C.c_int(0)
        _lib.check(self.lib.cib_engine_launch_count(self.engine, C.byref(t), C.byref(v)))
        return t.value, v.value
