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
    def __init__(self):
        super(IPyParallelLogger, self).__init__()
        self.history = {}

    def on_train_begin(self, logs=None):
        self.history = {"acc": [], "loss": [], "val_acc": [], "val_loss": [], "epoch": []}
        publish_data({"status": "Begin Training", "history": self.history})

    def on_train_end(self, logs=None):
        publish_data({"status": "Ended Training", "history": self.history})

    def on_epoch_begin(self, epoch, logs=None):
        if abort_requested():
            self.model.stop_training = True
        publish_data({"status": "Begin Epoch", "epoch": epoch, "history": self.history})

    def on_epoch_end(self, epoch, logs=None):
        for k in (logs or {}):
            self.history.setdefault(k, []).append(logs[k])
        self.history["epoch"].append(epoch)
        if abort_requested():
            self.model.stop_training = True
        publish_data({"status": "Ended Epoch", "epoch": epoch, "history": self.history})


class SessionConfig(object):
    def __init__(self, inter_op_parallelism_threads, intra_op_parallelism_threads):
        self.inter_op_parallelism_threads = inter_op_parallelism_threads
        self.intra_op_parallelism_threads = intra_op_parallelism_threads

    def __repr__(self):
        return "SessionConfig(inter_op=%d, intra_op=%d)" % (self.inter_op_parallelism_threads, self.intra_op_parallelism_threads)


def configure_session():
    """Thread settings from NUM_INTER_THREADS / NUM_INTRA_THREADS (defaults 2 / 32, as in the reference)."""
    n_inter_threads = int(os.environ.get('NUM_INTER_THREADS', 2))
    n_intra_threads = int(os.environ.get('NUM_INTRA_THREADS', 32))
    try:
        import torch
        torch.set_num_threads(max(1, min(n_intra_threads, os.cpu_count() or 1)))
    except Exception:
        pass
    return SessionConfig(n_inter_threads, n_intra_threads)
