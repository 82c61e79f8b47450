"""ATLAS RPV classifier: HDF5 loader, CNN builder (optional data-parallel wrap) and training driver.

Same public surface as the reference ``rpv.py``:

* ``load_file(filename, n_samples)`` -> data (n,64,64,1), labels, weights from HDF5 group ``all_events`` with
  datasets ``hist`` / ``y`` / ``weight`` (reference rpv.py:19-25), read with our own HDF5 reader (no h5py here).
* ``load_dataset(path, n_train, n_valid, n_test)`` -> three triples (rpv.py:27-36).  If the files are absent,
  deterministic synthetic RPV-shaped events are generated instead (there is no network on the build machines);
  set ``RPV_CHANNELS=3`` to get 64x64x3 images as BASELINE.json asks for the headline benchmark.
* ``build_model(input_shape, conv_sizes, fc_sizes, dropout, optimizer, lr, use_horovod)`` (rpv.py:38-72).
* ``train_model(...)`` (rpv.py:74-106) -- with a *fresh* callback list per call (the reference mutated its
  default argument).
"""
import os

import numpy as np

from .. import nn
from ..nn import layers, optimizers, callbacks as nn_callbacks
from ..utils import synthetic


def load_file(filename, n_samples):
    """First ``n_samples`` events of an ATLAS-RPV HDF5 file -> images (n,64,64,1), labels, event weights."""
    from ..utils import hdf5
    columns = {}
    with hdf5.File(filename, "r") as f:
        events = f["all_events"]
        for name in ("hist", "y", "weight"):
            columns[name] = np.asarray(events[name][:n_samples], np.float32)
    return columns["hist"][..., np.newaxis], columns["y"], columns["weight"]


_SPLITS = (("train.h5", 101), ("val.h5", 202), ("test.h5", 303))      # file name, seed of the synthetic stand-in


def _synthetic_split(n, seed):
    ch = int(os.environ.get("RPV_CHANNELS", "1"))
    if n <= 0:
        return (np.zeros((0, 64, 64, ch), np.float32), np.zeros((0,), np.float32), np.zeros((0,), np.float32))
    return synthetic.rpv_like(n, seed=seed, channels=ch)


_DATASET_CACHE = {}      # (path, counts, channels) -> splits; lives as long as the process (a cluster engine serves many trials)


def load_dataset(path, n_train=412416, n_valid=137471, n_test=137471):
    """(train, valid, test) triples of (images, labels, weights); synthetic events when the three files are not all there.

    The result is cached per process: a long-lived engine that runs one HPO trial after the other reads the files / generates the
    synthetic events once, not once per trial, and because the SAME arrays come back every time the GPU backend's device-resident
    copy (nn/backend_gpu.py, cached per host array) is reused across the trials' models too.  Treat the arrays as read-only."""
    counts = (n_train, n_valid, n_test)
    key = (os.path.abspath(path), counts, os.environ.get("RPV_CHANNELS", "1"))
    hit = _DATASET_CACHE.get(key)
    if hit is not None:
        return hit
    files = [os.path.join(path, name) for name, _ in _SPLITS]
    if all(os.path.exists(f) for f in files):
        splits = tuple(load_file(f, n) for f, n in zip(files, counts))
    else:
        splits = tuple(_synthetic_split(n, seed) for n, (_, seed) in zip(counts, _SPLITS))
    if len(_DATASET_CACHE) >= 4:
        _DATASET_CACHE.clear()
    _DATASET_CACHE[key] = splits
    return splits


def _rpv_stack(input_shape, conv_plan, fc_sizes, dropout):
    """Layer chain of the classifier as a functional graph: conv blocks, one dropout, flatten, dense blocks, sigmoid unit.
    ``conv_plan`` = [(filters, stride, pooled)]."""
    inputs = layers.Input(shape=input_shape)
    t = inputs
    for filters, stride, pooled in conv_plan:
        t = layers.Conv2D(int(filters), kernel_size=(3, 3), strides=(stride, stride), activation='relu', padding='same')(t)
        if pooled:
            t = layers.MaxPooling2D(pool_size=(2, 2))(t)
    if dropout is not None:
        t = layers.Dropout(dropout)(t)
    t = layers.Flatten()(t)
    for units in fc_sizes:
        t = layers.Dense(int(units), activation='relu')(t)
        if dropout is not None:
            t = layers.Dropout(dropout)(t)
    return inputs, layers.Dense(1, activation='sigmoid')(t)


def build_model(input_shape, conv_sizes=(8, 16, 32), fc_sizes=(64,), dropout=0.5, optimizer='Adam', lr=0.001,
                use_horovod=False, device='auto'):
    inputs, outputs = _rpv_stack(input_shape, [(c, 1, True) for c in conv_sizes], fc_sizes, dropout)
    opt = optimizers.get(optimizer, lr=lr)                    # library class by name, learning rate overridden (rpv.py:62)
    if use_horovod:
        from ..parallel import dist as hvd
        opt = hvd.DistributedOptimizer(opt)                   # gradients are all-reduce-averaged before the update
    model = nn.Model(inputs=inputs, outputs=outputs, name='RPVClassifier')
    model.compile(optimizer=opt, loss='binary_crossentropy', metrics=['accuracy'], device=device)
    return model


def build_legacy_model(input_shape, h1=64, h2=128, h3=256, h4=256, h5=512, optimizer='Adam', device='auto'):
    """The old single-GPU model of Train_rpv.ipynb:204-218,289-291: four 'same' convolutions with strides 1,2,1,2
    (no pooling) + Dense(h5).  34,515,201 parameters at the defaults on (64,64,1) (Train_rpv.ipynb:281)."""
    inputs, outputs = _rpv_stack(input_shape, [(h1, 1, False), (h2, 2, False), (h3, 1, False), (h4, 2, False)], [h5], None)
    model = nn.Model(inputs=inputs, outputs=outputs, name='RPVClassifier')
    model.compile(optimizer=optimizer, loss='binary_crossentropy', metrics=['accuracy'], device=device)
    return model


def train_model(model, train_input, train_labels, valid_input, valid_labels, batch_size, n_epochs,
                lr_warmup_epochs=0, lr_reduce_patience=8, checkpoint_file=None, use_horovod=False, verbose=2,
                callbacks=None):
    """Train the model; returns the ``History``.  A fresh callback list is assembled on every call."""
    stack = list(callbacks or [])
    if use_horovod:
        from ..parallel import dist as hvd
        stack.extend([hvd.callbacks.BroadcastGlobalVariablesCallback(0),          # same start on every rank
                      hvd.callbacks.MetricAverageCallback(),                      # same val_loss on every rank
                      hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=lr_warmup_epochs, verbose=1)])
    stack.append(nn_callbacks.ReduceLROnPlateau(patience=lr_reduce_patience, verbose=1))
    if checkpoint_file is not None:
        stack.append(nn_callbacks.ModelCheckpoint(checkpoint_file))
    return model.fit(x=train_input, y=train_labels, batch_size=batch_size, epochs=n_epochs,
                     validation_data=(valid_input, valid_labels), callbacks=stack, verbose=verbose)
