"""Callbacks with Keras 2.2 semantics: Callback protocol, History, ReduceLROnPlateau, ModelCheckpoint, EarlyStopping.

Reference: ReduceLROnPlateau(patience, verbose=1) at rpv.py:96-97; ModelCheckpoint(checkpoint_file) at rpv.py:100-101
and DistHPO_mnist.ipynb:184-185; the custom-callback protocol used by mlextras.py:8-33.
"""
import numpy as np


class Callback(object):
    """Hooks: on_train_begin/end(logs), on_epoch_begin/end(epoch, logs), on_batch_begin/end(batch, logs)."""

    def __init__(self):
        self.model = None
        self.params = {}

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = params

    def on_train_begin(self, logs=None):
        pass

    def on_train_end(self, logs=None):
        pass

    def on_epoch_begin(self, epoch, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass

    def on_batch_begin(self, batch, logs=None):
        pass

    def on_batch_end(self, batch, logs=None):
        pass


class CallbackList(object):
    def __init__(self, callbacks, model, params):
        self.callbacks = list(callbacks)
        for c in self.callbacks:
            c.set_model(model)
            c.set_params(params)

    def call(self, hook, *args):
        for c in self.callbacks:
            getattr(c, hook)(*args)


class History(Callback):
    """``history.history`` = dict of per-epoch lists, ``history.epoch`` = epoch indices (DistTrain_rpv.ipynb:310-311)."""

    def on_train_begin(self, logs=None):
        self.epoch = []
        self.history = {}

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


class ReduceLROnPlateau(Callback):
    """monitor=val_loss, factor=0.1, min_delta=1e-4, cooldown=0, min_lr=0 (Keras defaults); adds 'lr' to the logs."""

    def __init__(self, monitor="val_loss", factor=0.1, patience=10, verbose=0, mode="auto", min_delta=1e-4,
                 cooldown=0, min_lr=0.0):
        super(ReduceLROnPlateau, self).__init__()
        if factor >= 1.0:
            raise ValueError("ReduceLROnPlateau does not support a factor >= 1.0")
        self.monitor, self.factor, self.patience, self.verbose = monitor, float(factor), int(patience), verbose
        self.min_delta, self.cooldown, self.min_lr = float(min_delta), int(cooldown), float(min_lr)
        self.mode = "max" if (mode == "max" or (mode == "auto" and "acc" in monitor)) else "min"
        self._reset()

    def _reset(self):
        self.best = -np.inf if self.mode == "max" else np.inf
        self.wait = 0
        self.cooldown_counter = 0

    def _better(self, a, b):
        return a > b + self.min_delta if self.mode == "max" else a < b - self.min_delta

    def on_train_begin(self, logs=None):
        self._reset()

    def on_epoch_end(self, epoch, logs=None):
        logs = logs if logs is not None else {}
        logs["lr"] = float(self.model.get_lr())
        current = logs.get(self.monitor)
        if current is None:
            return
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.wait = 0
        if self._better(current, self.best):
            self.best = current
            self.wait = 0
        elif self.cooldown_counter <= 0:
            self.wait += 1
            if self.wait >= self.patience:
                old_lr = float(self.model.get_lr())
                if old_lr > self.min_lr:
                    new_lr = max(old_lr * self.factor, self.min_lr)
                    self.model.set_lr(new_lr)
                    if self.verbose > 0:
                        print("\nEpoch %05d: ReduceLROnPlateau reducing learning rate to %s." % (epoch + 1, new_lr))
                    self.cooldown_counter = self.cooldown
                    self.wait = 0


class ModelCheckpoint(Callback):
    """Saves the full model (architecture + weights + optimizer state) every epoch by default, overwriting ``filepath``
    unless it contains ``{epoch}``-style placeholders.  ``save_best_only`` keeps the best ``monitor`` instead.
    In data-parallel runs only rank 0 writes (the reference had no such guard)."""

    def __init__(self, filepath, monitor="val_loss", verbose=0, save_best_only=False, save_weights_only=False,
                 mode="auto", period=1):
        super(ModelCheckpoint, self).__init__()
        self.filepath, self.monitor, self.verbose = filepath, monitor, verbose
        self.save_best_only, self.save_weights_only, self.period = save_best_only, save_weights_only, int(period)
        self.mode = "max" if (mode == "max" or (mode == "auto" and ("acc" in monitor or monitor.startswith("fmeasure")))) else "min"
        self.best = -np.inf if self.mode == "max" else np.inf
        self.epochs_since_last_save = 0

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        self.epochs_since_last_save += 1
        if self.epochs_since_last_save < self.period:
            return
        self.epochs_since_last_save = 0
        from ..parallel import dist
        if dist.is_initialized() and dist.rank() != 0:
            return
        path = self.filepath.format(epoch=epoch + 1, **logs)
        if self.save_best_only:
            current = logs.get(self.monitor)
            if current is None:
                return
            better = current > self.best if self.mode == "max" else current < self.best
            if not better:
                return
            self.best = current
        if self.verbose > 0:
            print("\nEpoch %05d: saving model to %s" % (epoch + 1, path))
        if self.save_weights_only:
            self.model.save_weights(path)
        else:
            self.model.save(path)


class EarlyStopping(Callback):
    def __init__(self, monitor="val_loss", min_delta=0.0, patience=0, verbose=0, mode="auto"):
        super(EarlyStopping, self).__init__()
        self.monitor, self.min_delta, self.patience, self.verbose = monitor, abs(float(min_delta)), int(patience), verbose
        self.mode = "max" if (mode == "max" or (mode == "auto" and "acc" in monitor)) else "min"

    def on_train_begin(self, logs=None):
        self.wait = 0
        self.best = -np.inf if self.mode == "max" else np.inf
        self.stopped_epoch = 0

    def on_epoch_end(self, epoch, logs=None):
        current = (logs or {}).get(self.monitor)
        if current is None:
            return
        better = current - self.min_delta > self.best if self.mode == "max" else current + self.min_delta < self.best
        if better:
            self.best, self.wait = current, 0
        else:
            self.wait += 1
            if self.wait >= self.patience:
                self.stopped_epoch = epoch
                self.model.stop_training = True


class LambdaCallback(Callback):
    def __init__(self, on_epoch_begin=None, on_epoch_end=None, on_train_begin=None, on_train_end=None):
        super(LambdaCallback, self).__init__()
        for k, v in dict(on_epoch_begin=on_epoch_begin, on_epoch_end=on_epoch_end, on_train_begin=on_train_begin,
                         on_train_end=on_train_end).items():
            if v is not None:
                setattr(self, k, v)
