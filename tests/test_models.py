"""Model builders: parameter counts recorded in the reference notebooks are reproduced exactly."""
import io

import numpy as np

from cori_intml_examples_b200 import nn
from cori_intml_examples_b200.models import mnist, rpv
from cori_intml_examples_b200.nn import layers as L


def test_mnist_default_param_count():
    # GridSearchCV_mnist.ipynb:194 -> 37,562 parameters for build_model() defaults
    assert mnist.build_model(device="cpu").count_params() == 37562


def test_rpv_disttrain_param_count():
    # DistTrain_rpv.ipynb:257 -> 547,841 for conv [16,32,64], fc [128] on (64,64,1)
    m = rpv.build_model((64, 64, 1), conv_sizes=[16, 32, 64], fc_sizes=[128], dropout=0.2, device="cpu")
    assert m.count_params() == 547841
    assert m.name == "RPVClassifier"
    assert m.output_shape == (1,)


def test_rpv_legacy_param_count():
    # Train_rpv.ipynb:281 -> 34,515,201 for the old strided model
    assert rpv.build_legacy_model((64, 64, 1), device="cpu").count_params() == 34515201


def test_disttrain_mnist_inline_model_param_count():
    # DistTrain_mnist.ipynb:294-304 inline Sequential -> 1,199,882 parameters (:284)
    m = nn.Sequential()
    m.add(L.Conv2D(32, (3, 3), activation="relu", input_shape=(28, 28, 1)))
    m.add(L.Conv2D(64, (3, 3), activation="relu"))
    m.add(L.MaxPooling2D(pool_size=(2, 2)))
    m.add(L.Dropout(0.25))
    m.add(L.Flatten())
    m.add(L.Dense(128, activation="relu"))
    m.add(L.Dropout(0.5))
    m.add(L.Dense(10, activation="softmax"))
    assert m.count_params() == 1199882


def test_summary_format(fresh_names):
    m = mnist.build_model(device="cpu")
    buf = io.StringIO()
    m.summary(print_fn=lambda s: buf.write(s + "\n"))
    text = buf.getvalue()
    assert "Layer (type)                 Output Shape              Param #" in text
    assert "conv2d_1 (Conv2D)            (None, 26, 26, 4)         40" in text
    assert "Total params: 37,562" in text and "Non-trainable params: 0" in text


def test_shapes_and_layouts():
    m = rpv.build_model((64, 64, 3), conv_sizes=[8, 16, 32], fc_sizes=[64], device="cpu")
    ws = m.get_weights()
    assert ws[0].shape == (3, 3, 3, 8)          # HWIO
    assert ws[6].shape == (8 * 8 * 32, 64)      # Dense kernel (in, out), rows in (H, W, C) order
    x = np.random.RandomState(0).rand(5, 64, 64, 3).astype(np.float32)
    p = m.predict(x)
    assert p.shape == (5, 1) and np.all((p > 0) & (p < 1))
    assert m.predict_classes(x).shape == (5, 1)


def test_mnist_load_data_shapes():
    x_train, y_train, x_test, y_test = mnist.load_data(n_train=256, n_test=64)
    assert x_train.shape == (256, 28, 28, 1) and x_train.dtype == np.float32
    assert y_train.shape == (256, 10) and np.allclose(y_train.sum(1), 1)
    assert 0.0 <= x_train.min() and x_train.max() <= 1.0 and y_test.shape == (64, 10)


def test_rpv_load_dataset_synthetic_and_hdf5(tmp_path):
    (tx, ty, tw), (vx, vy, vw), (sx, sy, sw) = rpv.load_dataset(str(tmp_path / "missing"), 32, 16, 8)
    assert tx.shape == (32, 64, 64, 1) and vy.shape == (16,) and sw.shape == (8,)
    # real-file path: write the reference's HDF5 layout (group all_events: hist, y, weight) and read it back
    from cori_intml_examples_b200.utils import hdf5
    for name, n in (("train", 12), ("val", 6), ("test", 4)):
        with hdf5.File(str(tmp_path / (name + ".h5")), "w") as f:
            g = f.create_group("all_events")
            g.create_dataset("hist", data=np.random.rand(n, 64, 64).astype(np.float32))
            g.create_dataset("y", data=(np.arange(n) % 2).astype(np.float32))
            g.create_dataset("weight", data=np.ones(n, np.float32))
    (tx, ty, tw), _, _ = rpv.load_dataset(str(tmp_path), 10, 6, 4)
    assert tx.shape == (10, 64, 64, 1) and ty.tolist() == [0, 1] * 5
