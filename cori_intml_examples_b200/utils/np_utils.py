"""Small array helpers (Keras ``to_categorical``, reference mnist.py:40-41)."""
import numpy as np


def to_categorical(y, num_classes=None):
    y = np.asarray(y, dtype="int64").ravel()
    if num_classes is None:
        num_classes = int(y.max()) + 1
    out = np.zeros((y.shape[0], num_classes), dtype=np.float32)
    out[np.arange(y.shape[0]), y] = 1.0
    return out
