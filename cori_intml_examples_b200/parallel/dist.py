"""Horovod-style data-parallel façade: init/rank/local_rank/size, DistributedOptimizer and the three callbacks.

One process per GPU.  The *control* exchanges (rendezvous, IPC-handle exchange, CPU-oracle collectives) go through
``torch.distributed`` with the **gloo** backend; the *data plane* on GPUs is the hand-written peer-to-peer kernel
(csrc/allreduce.cuh) working on CUDA-IPC-mapped arenas -- NCCL is not on this path.  On machines without a GPU
(the CPU test tier) gradients are averaged with gloo all-reduce, which gives the same semantics.

Reference: hvd.init/rank/local_rank/size (train_rpv.py:37-39), hvd.DistributedOptimizer (rpv.py:63-65),
hvd.callbacks.* (rpv.py:83-93).  Like the reference, data is NOT sharded: every rank trains on the same first-n
samples with its own shuffle (DistTrain_mnist.ipynb:224), so global batch = size x batch_size.
"""
import ctypes as C
import os

import numpy as np

from ..nn import callbacks as _cb

_state = {"initialized": False, "rank": 0, "size": 1, "local_rank": 0, "own_pg": False}


def init(rank=None, size=None, master_addr=None, master_port=None, local_rank=None, persistent=False):
    """Bootstrap the communicator inside already-running engine processes (or under torchrun).

    With no arguments the torchrun environment (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT) is used;
    with nothing set this is a single-rank world.  ``persistent=True`` (long-lived cluster engines that serve many training jobs as
    one data-parallel group, hpo.Evaluator(launcher="cluster")): later ``shutdown()`` calls made by those jobs leave the group up;
    ``shutdown(force=True)`` ends it."""
    import torch.distributed as td
    if persistent:
        _state["persistent"] = True
    if _state["initialized"]:
        return
    rank = int(os.environ.get("RANK", 0)) if rank is None else int(rank)
    size = int(os.environ.get("WORLD_SIZE", 1)) if size is None else int(size)
    if local_rank is None:
        # torchrun exports LOCAL_RANK; a cluster engine that shares GPU visibility is told its GPU through CIB_DEVICE
        local_rank = os.environ.get("LOCAL_RANK", os.environ.get("CIB_DEVICE", rank))
    local_rank = int(local_rank)
    if size > 1 and not td.is_initialized():
        addr = master_addr or os.environ.get("MASTER_ADDR", "127.0.0.1")
        port = int(master_port or os.environ.get("MASTER_PORT", 29511))
        td.init_process_group("gloo", init_method="tcp://%s:%d" % (addr, port), rank=rank, world_size=size)
        _state["own_pg"] = True
    _state.update(initialized=True, rank=rank, size=size, local_rank=local_rank)
    try:
        import torch
        if torch.cuda.is_available():
            torch.cuda.set_device(local_rank % torch.cuda.device_count())
    except Exception:
        pass


def shutdown(force=False):
    import torch.distributed as td
    if _state.get("persistent") and not force:
        return
    _state["persistent"] = False
    if _state["own_pg"] and td.is_initialized():
        td.destroy_process_group()
    _state.update(initialized=False, rank=0, size=1, local_rank=0, own_pg=False, symbuf=None)


def is_initialized():
    return _state["initialized"]


def rank():
    return _state["rank"]


def local_rank():
    return _state["local_rank"]


def size():
    return _state["size"]


def _group():
    import torch.distributed as td
    return td.group.WORLD if td.is_initialized() else None


def _p2p_allreduce(array, average):
    """Small all-reduce on the NVLink fabric (one-shot peer-to-peer kernel on an IPC-mapped 4 KB buffer set up next to the
    engine communicator) -- SURVEY 2.7.2 C3: the per-epoch metric average.  Returns None when that path is not available."""
    sb = _state.get("symbuf")
    if sb is None or np.size(array) > 1024 or os.environ.get("CIB_P2P_CONTROL", "1") == "0":
        return None
    import torch
    lib, handle, view = sb
    n = int(np.size(array))
    npad = (n + 3) // 4 * 4
    host = np.zeros(npad, np.float32)
    host[:n] = np.asarray(array, np.float64).reshape(-1)
    view[:npad].copy_(torch.from_numpy(host))
    from ..ops import _lib
    _lib.check(lib.cib_symbuf_allreduce(handle, npad, 1 if average else 0, 1, None))
    out = view[:n].cpu().numpy().astype(np.float64).reshape(np.shape(array))
    return out if np.ndim(array) else float(out)


def allreduce(array, average=True):
    """All-reduce a NumPy array / scalar over the ranks: on the peer-to-peer fabric once a GPU communicator exists (small
    arrays: metrics), otherwise through gloo on the host (CPU oracle, bootstrap)."""
    import torch
    import torch.distributed as td
    if size() == 1:
        return array
    fast = _p2p_allreduce(array, average)
    if fast is not None:
        return fast
    t = torch.from_numpy(np.array(array, dtype=np.float64, copy=True).reshape(-1))
    td.all_reduce(t)
    if average:
        t /= size()
    out = t.numpy().reshape(np.shape(array))
    return out if np.ndim(array) else float(out)


def broadcast(array, root=0):
    import torch
    import torch.distributed as td
    if size() == 1:
        return array
    t = torch.from_numpy(np.ascontiguousarray(array).copy())
    td.broadcast(t, src=root)
    return t.numpy()


def shard(*arrays):
    """This rank's 1/size slice of each array (contiguous block, remainder dropped so that every rank runs the same number of
    steps).  The reference does NOT shard -- every rank trains on all samples with its own shuffle (DistTrain_mnist.ipynb:224,
    SURVEY 2.4.2) -- so this is the optional "sharded-data" variant of SURVEY 7.5; throughput numbers must say which one they use."""
    n = min(len(a) for a in arrays)
    per = n // size()
    lo = rank() * per
    out = tuple(a[lo:lo + per] for a in arrays)
    return out[0] if len(out) == 1 else out


def allgather_object(obj):
    import torch.distributed as td
    if size() == 1:
        return [obj]
    out = [None] * size()
    td.all_gather_object(out, obj)
    return out


def barrier():
    import torch.distributed as td
    if size() > 1:
        td.barrier()


def _setup_nvls(lib, engine):
    """Bind the engine's gradient buffer into one NVSwitch multicast object shared by all ranks (csrc/nvls_drv.cuh).  Collective.
    Rank 0 creates the object and exports it as a POSIX file descriptor, which travels to the other ranks over a Unix-domain socket
    (SCM_RIGHTS); every step is followed by an agreement round, so either all ranks end up on the NVLS path or none does."""
    import socket
    import tempfile
    import time
    import uuid
    W, R = size(), rank()

    def agree(ok):
        if not ok and os.environ.get("CIB_NVLS_DEBUG"):
            import sys
            sys.stderr.write("[nvls rank %d] %s\n" % (R, lib.cib_last_error().decode("utf8", "replace")))
        return all(allgather_object(bool(ok)))

    fd = C.c_int(-1)

    def undo():
        lib.cib_engine_nvls_step(engine, 9, W, -1, None)
        if R == 0 and fd.value >= 0:
            try:
                os.close(fd.value)
            except OSError:
                pass
            fd.value = -1
        return False
    if not agree(lib.cib_engine_nvls_step(engine, 0, W, -1, None) == 0):
        return undo()
    srv, path, ok = None, "", True
    if R == 0:
        ok = lib.cib_engine_nvls_step(engine, 1, W, -1, C.byref(fd)) == 0
        if ok:
            path = os.path.join(tempfile.gettempdir(), "cib_nvls_%s.sock" % uuid.uuid4().hex)
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(path); srv.listen(W)
    path = allgather_object(path)[0]
    if not path:
        agree(False)
        return undo()
    try:
        if R == 0:
            srv.settimeout(60)
            for _ in range(W - 1):
                conn, _ = srv.accept()
                socket.send_fds(conn, [b"mc"], [fd.value])
                conn.recv(1)                                  # the peer has the descriptor
                conn.close()
            srv.close(); os.unlink(path); os.close(fd.value); fd.value = -1
        else:
            cl = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            deadline = time.time() + 60
            while True:
                try:
                    cl.connect(path); break
                except OSError:
                    if time.time() > deadline:
                        raise
                    time.sleep(0.01)
            _, fds, _, _ = socket.recv_fds(cl, 16, 1)
            cl.send(b"k"); cl.close()
            ok = lib.cib_engine_nvls_step(engine, 1, W, int(fds[0]), None) == 0
            os.close(fds[0])
    except Exception:
        ok = False
    if not agree(ok):
        return undo()
    if not agree(lib.cib_engine_nvls_step(engine, 2, W, -1, None) == 0):      # every device is in the group ...
        return undo()
    if not agree(lib.cib_engine_nvls_step(engine, 3, W, -1, None) == 0):      # ... before anybody binds memory
        return undo()
    return True


class DistributedOptimizer(object):
    """Wraps an optimizer: every gradient is all-reduce-averaged across ranks before the update.

    GPU backend: the engine's optimizer node is replaced by ONE fused kernel doing barrier -> peer reduce ->
    1/N scale -> optimizer update -> parameter all-gather (csrc/allreduce.cuh).  CPU backend: gloo all-reduce of
    the flattened gradient list ("fusion buffer"), then the NumPy optimizer."""

    def __init__(self, optimizer, mode=0):
        from ..nn import optimizers
        self.optimizer = optimizers.get(optimizer)
        self.mode = mode or int(os.environ.get("CIB_AR_MODE", "0"))     # 0 auto, 1 one-shot, 2 two-shot
        self.model = None
        self._comm = None

    def bind(self, model):
        self.model = model

    # ---- CPU oracle path
    def allreduce_gradients(self, grads):
        if size() == 1:
            return grads
        import torch
        import torch.distributed as td
        flat = torch.from_numpy(np.concatenate([g.reshape(-1) for g in grads]).astype(np.float32))
        td.all_reduce(flat)
        flat /= size()
        out, off = [], 0
        arr = flat.numpy()
        for g in grads:
            out.append(arr[off:off + g.size].reshape(g.shape)); off += g.size
        return out

    # ---- GPU path: called by GpuBackend once its engine exists
    def setup_engine(self, backend):
        if size() == 1:
            return
        from ..ops import _lib
        lib = backend.lib
        if self._comm is not None:                      # the engine it belonged to is gone (batch size changed): release the mappings
            lib.cib_comm_destroy(self._comm)
            self._comm = None
        handle = (C.c_ubyte * 64)()
        _lib.check(lib.cib_engine_ipc_handle(backend.engine, handle))
        handles = allgather_object(bytes(handle))
        blob = (C.c_ubyte * (64 * size())).from_buffer_copy(b"".join(handles))
        comm = C.c_void_p()
        _lib.check(lib.cib_comm_create_ipc(C.byref(comm), backend.engine, rank(), size(), blob))
        self._comm = comm
        _lib.check(lib.cib_engine_attach_comm(backend.engine, comm, rank(), self.mode))
        backend._dist_comm_attached = True
        # in-switch reduction (NVLS) for the large gradient bucket where the platform exposes multicast: CIB_NVLS=1
        backend._nvls = bool(os.environ.get("CIB_NVLS", "0") == "1" and _setup_nvls(lib, backend.engine))
        if _state.get("symbuf") is None and os.environ.get("CIB_P2P_CONTROL", "1") != "0":
            # 4 KB symmetric buffer for the small control collectives (metric averages)
            sb = C.c_void_p()
            _lib.check(lib.cib_symbuf_create(C.byref(sb), backend.device, 1024))
            h = (C.c_ubyte * 64)()
            _lib.check(lib.cib_symbuf_ipc_handle(sb, h))
            hs = allgather_object(bytes(h))
            allh = (C.c_ubyte * (64 * size())).from_buffer_copy(b"".join(hs))
            _lib.check(lib.cib_symbuf_attach_ipc(sb, rank(), size(), allh))
            p = C.c_void_p()
            _lib.check(lib.cib_symbuf_ptr(sb, C.byref(p)))
            import torch

            class _View(object):
                pass
            v = _View()
            v.__cuda_array_interface__ = {"shape": (1024,), "typestr": "<f4", "data": (p.value, False), "version": 3}
            with torch.cuda.device(backend.device):
                _state["symbuf"] = (lib, sb, torch.as_tensor(v, device="cuda:%d" % backend.device))
        barrier()


class _Callbacks(object):
    class BroadcastGlobalVariablesCallback(_cb.Callback):
        """At train begin, broadcast all model AND optimizer variables from ``root_rank`` (rpv.py:85)."""

        def __init__(self, root_rank=0):
            super(_Callbacks.BroadcastGlobalVariablesCallback, self).__init__()
            self.root_rank = root_rank

        def on_train_begin(self, logs=None):
            if size() == 1:
                return
            m = self.model
            be = m.backend
            if hasattr(be, "_ensure") and os.environ.get("CIB_P2P_CONTROL", "1") != "0":
                be._ensure(int((getattr(self, "params", None) or {}).get("batch_size") or be.batch or 32))   # creates the engine + communicator
            if getattr(be, "_dist_comm_attached", False) and be.engine is not None and os.environ.get("CIB_P2P_CONTROL", "1") != "0":
                # GPU communicator present: every rank pulls the root's parameter and slot buffers over NVLink (C2)
                from ..ops import _lib
                m.optimizer.iterations = int(broadcast(np.array([m.optimizer.iterations], np.int64), self.root_rank)[0])
                if hasattr(m.optimizer, "m_schedule"):
                    m.optimizer.m_schedule = float(broadcast(np.array([m.optimizer.m_schedule], np.float64), self.root_rank)[0])
                _lib.check(be.lib.cib_engine_gather_slots(be.engine))      # the root's moments may be sharded from an earlier fit()
                _lib.check(be.lib.cib_engine_broadcast_state(be.engine, int(self.root_rank)))
                _lib.check(be.lib.cib_engine_sync(be.engine))
                be.push_hyper()
                barrier()
                return
            ws = m.backend.get_weights()
            ws = [broadcast(w, self.root_rank) for w in ws]
            m.backend.set_weights(ws)
            slots = m.backend.get_slots()
            slots = [[broadcast(a, self.root_rank) for a in s] for s in slots]
            m.backend.set_slots(slots)
            m.optimizer.iterations = int(broadcast(np.array([m.optimizer.iterations], np.int64), self.root_rank)[0])
            if hasattr(m.backend, "push_hyper"):
                m.backend.push_hyper()

    class MetricAverageCallback(_cb.Callback):
        """At epoch end, all-reduce-average every scalar in ``logs`` so all ranks take identical decisions (rpv.py:87)."""

        def on_epoch_end(self, epoch, logs=None):
            if size() == 1 or not logs:
                return
            keys = sorted(k for k, v in logs.items() if np.isscalar(v))
            vals = allreduce(np.array([logs[k] for k in keys], np.float64), average=True)
            for k, v in zip(keys, vals):
                logs[k] = float(v)

    class LearningRateWarmupCallback(_cb.Callback):
        """Goyal et al. warm-up from lr/size to lr over ``warmup_epochs`` (rpv.py:91-92; the reference passes 0 = no-op)."""

        def __init__(self, warmup_epochs=5, momentum_correction=True, steps_per_epoch=None, verbose=0):
            super(_Callbacks.LearningRateWarmupCallback, self).__init__()
            self.warmup_epochs, self.verbose = warmup_epochs, verbose
            self.initial_lr = None

        def on_train_begin(self, logs=None):
            self.initial_lr = self.model.get_lr()

        def on_epoch_begin(self, epoch, logs=None):
            if self.warmup_epochs <= 0 or epoch >= self.warmup_epochs or size() == 1:
                return
            n = size()
            lr = self.initial_lr * (1.0 / n) * (epoch * (n - 1) / float(self.warmup_epochs) + 1)
            self.model.set_lr(lr)

        def on_epoch_end(self, epoch, logs=None):
            if self.warmup_epochs > 0 and epoch == self.warmup_epochs - 1:
                self.model.set_lr(self.initial_lr)
                if self.verbose and rank() == 0:
                    print("\nEpoch %d: finished gradual learning rate warmup to %g." % (epoch + 1, self.initial_lr))


callbacks = _Callbacks
