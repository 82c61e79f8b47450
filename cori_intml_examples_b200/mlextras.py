"""``IPyParallelLogger`` callback and ``configure_session`` -- same names and behaviour as the reference ``mlextras.py``.

``IPyParallelLogger`` publishes ``{"status", ["epoch"], "history"}`` to the client through ``publish_data`` at train
begin/end and epoch begin/end (reference mlextras.py:8-33; status strings "Begin Training", "Ended Training",
"Begin Epoch", "Ended Epoch").  It also turns a client-side abort into ``model.stop_training`` between epochs, which
is what gives the HPO dashboard a working Stop button.

``configure_session`` configured TensorFlow thread pools on Cori (mlextras.py:35-43).  The GPU analogue is the
per-trial execution resources; the returned ``SessionConfig`` records the requested values and applies the CPU
thread count to PyTorch (used by the CPU oracle backend).
"""
import os

from .nn.callbacks import Callback
from .cluster.datapub import publish_data, abort_requested


class IPyParallelLogger(Callback):
    KEYS = ("acc", "loss", "val_acc", "val_loss", "epoch")

    def __init__(self):
        super(IPyParallelLogger, self).__init__()
        self.history = {}

    def _publish(self, status, epoch=None):
        msg = {"status": status, "history": self.history}
        if epoch is not None:
            msg["epoch"] = epoch
        publish_data(msg)

    def _honour_abort(self):
        if abort_requested():                       # Stop button / AsyncResult.abort(): finish the epoch, then stop
            self.model.stop_training = True

    def on_train_begin(self, logs=None):
        self.history = {k: [] for k in self.KEYS}
        self._publish("Begin Training")

    def on_train_end(self, logs=None):
        self._publish("Ended Training")

    def on_epoch_begin(self, epoch, logs=None):
        self._honour_abort()
        self._publish("Begin Epoch", epoch)

    def on_epoch_end(self, epoch, logs=None):
        for name, value in (logs or {}).items():
            self.history.setdefault(name, []).append(value)
        self.history["epoch"].append(epoch)
        self._honour_abort()
        self._publish("Ended Epoch", epoch)


class SessionConfig(object):
    def __init__(self, inter_op_parallelism_threads, intra_op_parallelism_threads):
        self.inter_op_parallelism_threads = inter_op_parallelism_threads
        self.intra_op_parallelism_threads = intra_op_parallelism_threads

    def __repr__(self):
        return "SessionConfig(inter_op=%d, intra_op=%d)" % (self.inter_op_parallelism_threads, self.intra_op_parallelism_threads)


def configure_session():
    """Thread settings from NUM_INTER_THREADS / NUM_INTRA_THREADS (defaults 2 / 32, as in the reference)."""
    inter, intra = (int(os.environ.get(name, default)) for name, default in (("NUM_INTER_THREADS", 2), ("NUM_INTRA_THREADS", 32)))
    try:
        import torch
        torch.set_num_threads(max(1, min(intra, os.cpu_count() or 1)))
    except Exception:
        pass
    return SessionConfig(inter, intra)
