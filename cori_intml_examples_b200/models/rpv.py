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
    from ..utils import hdf5
    with hdf5.File(filename, "r") as f:
        g = f["all_events"]
        data = np.asarray(g["hist"][:n_samples], np.float32)[:, :, :, None]
        labels = np.asarray(g["y"][:n_samples], np.float32)
        weights = np.asarray(g["weight"][:n_samples], np.float32)
    return data, labels, weights


def _synthetic_split(n, seed):
    ch = int(os.environ.get("RPV_CHANNELS", "1"))
    if n <= 0:
        return (np.zeros((0, 64, 64, ch), np.float32), np.zeros((0,), np.float32), np.zeros((0,), np.float32))
    return synthetic.rpv_like(n, seed=seed, channels=ch)


def load_dataset(path, n_train=412416, n_valid=137471, n_test=137471):
    train_file = os.path.join(path, 'train.h5')
    valid_file = os.path.join(path, 'val.h5')
    test_file = os.path.join(path, 'test.h5')
    if all(os.path.exists(f) for f in (train_file, valid_file, test_file)):
        return (load_file(train_file, n_train), load_file(valid_file, n_valid), load_file(test_file, n_test))
    return (_synthetic_split(n_train, 101), _synthetic_split(n_valid, 202), _synthetic_split(n_test, 303))


def build_model(input_shape, conv_sizes=(8, 16, 32), fc_sizes=(64,), dropout=0.5, optimizer='Adam', lr=0.001,
                use_horovod=False, device='auto'):
    inputs = layers.Input(shape=input_shape)
    h = inputs
    conv_args = dict(kernel_size=(3, 3), activation='relu', padding='same')
    for conv_size in conv_sizes:
        h = layers.Conv2D(int(conv_size), **conv_args)(h)
        h = layers.MaxPooling2D(pool_size=(2, 2))(h)
    h = layers.Dropout(dropout)(h)
    h = layers.Flatten()(h)
    for fc_size in fc_sizes:
        h = layers.Dense(int(fc_size), activation='relu')(h)
        h = layers.Dropout(dropout)(h)
    outputs = layers.Dense(1, activation='sigmoid')(h)

    opt = optimizers.get(optimizer, lr=lr)
    if use_horovod:
        from ..parallel import dist as hvd
        opt = hvd.DistributedOptimizer(opt)

    model = nn.Model(inputs=inputs, outputs=outputs, name='RPVClassifier')
    model.compile(optimizer=opt, loss='binary_crossentropy', metrics=['accuracy'], device=device)
    return model


def build_legacy_model(input_shape, h1=64, h2=128, h3=256, h4=256, h5=512, optimizer='Adam', device='auto'):
    """The old single-GPU model of Train_rpv.ipynb:204-218,289-291: four 'same' convolutions with strides 1,2,1,2
    (no pooling) + Dense(h5).  34,515,201 parameters at the defaults on (64,64,1) (Train_rpv.ipynb:281)."""
    inputs = layers.Input(shape=input_shape)
    h = inputs
    for c, s in ((h1, 1), (h2, 2), (h3, 1), (h4, 2)):
        h = layers.Conv2D(int(c), kernel_size=(3, 3), strides=(s, s), activation='relu', padding='same')(h)
    h = layers.Flatten()(h)
    h = layers.Dense(int(h5), activation='relu')(h)
    outputs = layers.Dense(1, activation='sigmoid')(h)
    model = nn.Model(inputs=inputs, outputs=outputs, name='RPVClassifier')
    model.compile(optimizer=optimizer, loss='binary_crossentropy', metrics=['accuracy'], device=device)
    return model


def train_model(model, train_input, train_labels, valid_input, valid_labels, batch_size, n_epochs,
                lr_warmup_epochs=0, lr_reduce_patience=8, checkpoint_file=None, use_horovod=False, verbose=2,
                callbacks=None):
    """Train the model; returns the ``History``."""
    cbs = list(callbacks) if callbacks else []
    if use_horovod:
        from ..parallel import dist as hvd
        cbs += [
            hvd.callbacks.BroadcastGlobalVariablesCallback(0),
            hvd.callbacks.MetricAverageCallback(),
            hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=lr_warmup_epochs, verbose=1),
        ]
    cbs.append(nn_callbacks.ReduceLROnPlateau(patience=lr_reduce_patience, verbose=1))
    if checkpoint_file is not None:
        cbs.append(nn_callbacks.ModelCheckpoint(checkpoint_file))
    return model.fit(x=train_input, y=train_labels, batch_size=batch_size, epochs=n_epochs,
                     validation_data=(valid_input, valid_labels), callbacks=cbs, verbose=verbose)
