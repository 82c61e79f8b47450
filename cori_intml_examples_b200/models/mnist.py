"""MNIST data loader + CNN builder -- same public surface as the reference ``mnist.py``.

``load_data()`` returns ``x_train, y_train, x_test, y_test`` with shapes (N,28,28,1) float32 in [0,1] and one-hot
(N,10) labels (reference mnist.py:32-42).  There is no network here, so unless ``MNIST_NPZ`` points at a local
``mnist.npz`` the data is the deterministic synthetic MNIST-shaped set (60000/10000 by default).
``build_model`` reproduces mnist.py:44-59 layer for layer (37,562 parameters at the defaults with h1=8... see tests).
"""
import os

import numpy as np

from ..nn import Sequential
from ..nn.layers import Conv2D, MaxPooling2D, Dropout, Flatten, Dense
from ..utils.np_utils import to_categorical
from ..utils import synthetic

n_classes = 10
img_rows, img_cols = 28, 28


def _images(raw):
    return raw.reshape(-1, img_rows, img_cols, 1).astype(np.float32) / 255


def load_data(n_train=60000, n_test=10000, seed=0):
    npz = os.environ.get("MNIST_NPZ")
    if npz and os.path.exists(npz):
        with np.load(npz) as f:
            (x_train, y_train), (x_test, y_test) = (_images(f["x_train"]), f["y_train"]), (_images(f["x_test"]), f["y_test"])
    else:
        (x_train, y_train), (x_test, y_test) = synthetic.mnist_like(n_train, seed=seed), synthetic.mnist_like(n_test, seed=seed + 1)
    return x_train, to_categorical(y_train, n_classes), x_test, to_categorical(y_test, n_classes)


def build_model(h1=4, h2=8, h3=32, dropout=0.5, optimizer='Adadelta', device='auto'):
    """Conv(h1,3x3,relu) - Conv(h2,3x3,relu) - MaxPool(2) - Dropout - Flatten - Dense(h3,relu) - Dropout - Dense(10,softmax)."""
    chain = [Conv2D(int(h1), (3, 3), activation='relu', input_shape=(img_rows, img_cols, 1)),
             Conv2D(int(h2), (3, 3), activation='relu'),
             MaxPooling2D(pool_size=(2, 2)),
             Dropout(dropout),
             Flatten(),
             Dense(int(h3), activation='relu'),
             Dropout(dropout),
             Dense(n_classes, activation='softmax')]
    model = Sequential()
    for layer in chain:
        model.add(layer)
    model.compile(optimizer=optimizer, loss='categorical_crossentropy', metrics=['accuracy'], device=device)
    return model
