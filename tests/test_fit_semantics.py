"""Keras fit()/callback semantics the reference relies on."""
import numpy as np
import pytest

from cori_intml_examples_b200 import nn
from cori_intml_examples_b200.models import mnist
from cori_intml_examples_b200.nn import callbacks as C
from cori_intml_examples_b200.nn import layers as L
from cori_intml_examples_b200.utils import synthetic
from cori_intml_examples_b200.utils.np_utils import to_categorical


def tiny_mnist(n=600, seed=0):
    x, y = synthetic.mnist_like(n, seed)
    return x, to_categorical(y, 10)


def test_validation_split_takes_last_fraction_and_history_keys(capsys):
    x, y = tiny_mnist(600)
    m = mnist.build_model(4, 4, 8, 0.1, "Adam", device="cpu")
    h = m.fit(x, y, batch_size=128, epochs=2, validation_split=0.17, verbose=2)
    out = capsys.readouterr().out
    # int(600 * 0.83) = 498 train / 102 validation -- the same arithmetic gives 49800 / 10200 (DistHPO_mnist.ipynb:293)
    assert "Train on 498 samples, validate on 102 samples" in out
    assert int(60000 * (1 - 0.17)) == 49800
    assert "Epoch 1/2" in out and " - loss: " in out and " - val_acc: " in out
    assert sorted(h.history) == ["acc", "loss", "val_acc", "val_loss"]
    assert h.epoch == [0, 1] and all(len(v) == 2 for v in h.history.values())


def test_partial_last_batch_and_running_mean_weighting():
    x, y = tiny_mnist(200)
    m = mnist.build_model(4, 4, 8, 0.0, O_SGD(), device="cpu")
    seen = []

    class Spy(C.Callback):
        def on_epoch_end(self, epoch, logs=None):
            seen.append(dict(logs))
    m.fit(x, y, batch_size=64, epochs=1, verbose=0, shuffle=False, callbacks=[Spy()])
    # lr = 0: weights never move, so the epoch loss must equal the sample-weighted mean over 64,64,64,8-sized batches
    l, a = m.evaluate(x, y, batch_size=200, verbose=0)
    assert abs(seen[0]["loss"] - l) < 1e-5 and abs(seen[0]["acc"] - a) < 1e-6


def O_SGD():
    from cori_intml_examples_b200.nn import optimizers
    return optimizers.SGD(lr=0.0)


def test_reduce_lr_on_plateau_state_machine():
    cb = C.ReduceLROnPlateau(patience=2, verbose=0)

    class FakeModel(object):
        lr = 0.01
        def get_lr(self): return self.lr
        def set_lr(self, v): self.lr = v
    fm = FakeModel(); cb.set_model(fm); cb.on_train_begin()
    logs = [{"val_loss": v} for v in (1.0, 0.9, 0.95, 0.93, 0.92, 0.91)]
    lrs = []
    for e, lg in enumerate(logs):
        cb.on_epoch_end(e, lg); lrs.append(fm.lr); assert "lr" in lg
    # best=0.9 at epoch 1; epochs 2,3 do not improve by 1e-4 -> reduce at epoch 3; wait restarts; epochs 4,5 -> reduce again
    assert lrs[:3] == [0.01, 0.01, 0.01] and lrs[3] == pytest.approx(0.001) and lrs[5] == pytest.approx(0.0001)


def test_model_checkpoint_overwrites_and_best_only(tmp_path):
    x, y = tiny_mnist(256)
    m = mnist.build_model(4, 4, 8, 0.1, "Adam", device="cpu")
    path = str(tmp_path / "model.h5")
    m.fit(x, y, batch_size=64, epochs=2, validation_split=0.25, verbose=0, callbacks=[C.ModelCheckpoint(path)])
    m2 = nn.load_model(path, device="cpu")
    # default ModelCheckpoint = last epoch's weights
    for a, b in zip(m.get_weights(), m2.get_weights()):
        np.testing.assert_array_equal(a, b)
    best = C.ModelCheckpoint(str(tmp_path / "best.h5"), save_best_only=True)
    m.fit(x, y, batch_size=64, epochs=2, validation_split=0.25, verbose=0, callbacks=[best])
    assert np.isfinite(best.best)


def test_callbacks_are_fresh_per_train_model_call():
    from cori_intml_examples_b200.models import rpv
    x, y, _ = synthetic.rpv_like(48, 1)
    m = rpv.build_model((64, 64, 1), [4, 4, 8], [8], dropout=0.1, device="cpu")
    mine = []
    rpv.train_model(m, x[:32], y[:32], x[32:], y[32:], batch_size=16, n_epochs=1, verbose=0, callbacks=mine)
    rpv.train_model(m, x[:32], y[:32], x[32:], y[32:], batch_size=16, n_epochs=1, verbose=0)
    assert mine == []          # the reference mutated its default list (rpv.py:79); we must not


def test_extreme_dropout_does_not_produce_nans():
    x, y = tiny_mnist(256)
    m = mnist.build_model(8, 8, 16, 0.98, "Adadelta", device="cpu")      # DistHPO_mnist.ipynb trial 2: dropout 0.980
    h = m.fit(x, y, batch_size=64, epochs=1, validation_split=0.25, verbose=0)
    assert np.isfinite(h.history["loss"]).all() and np.isfinite(h.history["val_loss"]).all()


def test_training_learns_separable_data():
    np.random.seed(1)                       # the initialisers draw from NumPy's global stream: do not depend on the tests run before
    x, y = tiny_mnist(1024)
    m = mnist.build_model(8, 16, 32, 0.1, "Adam", device="cpu")
    h = m.fit(x, y, batch_size=64, epochs=3, validation_split=0.2, verbose=0)
    assert h.history["loss"][-1] < 0.5 * h.history["loss"][0] and h.history["val_acc"][-1] > 0.9


def test_keras_losses_and_accuracy():
    import torch
    from cori_intml_examples_b200.nn.backend_cpu import keras_loss_and_acc
    p = torch.tensor([[0.0], [1.0], [0.25]]); y = torch.tensor([[0.0], [0.0], [1.0]])
    loss, acc = keras_loss_and_acc(p, y, "binary_crossentropy")
    eps = np.float32(1e-7)      # Keras clips in float32 as well
    np.testing.assert_allclose(loss.numpy(), [-np.log(np.float32(1) - eps), -np.log(np.float32(1) - (np.float32(1) - eps)), -np.log(np.float32(0.25))], rtol=1e-4)
    assert acc.tolist() == [1.0, 0.0, 0.0]
    p = torch.tensor([[0.7, 0.2, 0.1], [0.0, 0.0, 1.0]]); y = torch.tensor([[1.0, 0, 0], [0, 1.0, 0]])
    loss, acc = keras_loss_and_acc(p, y, "categorical_crossentropy")
    np.testing.assert_allclose(loss.numpy(), [-np.log(np.float32(0.7)), -np.log(eps)], rtol=1e-4)
    assert acc.tolist() == [1.0, 0.0]


def test_functional_and_sequential_agree():
    rng = np.random.RandomState(0)
    x = rng.rand(4, 12, 12, 2).astype(np.float32)
    inp = L.Input((12, 12, 2)); h = L.Conv2D(4, (3, 3), activation="relu", padding="same")(inp)
    h = L.MaxPooling2D((2, 2))(h); h = L.Flatten()(h); h = L.Dense(5, activation="relu")(h); out = L.Dense(1, activation="sigmoid")(h)
    f = nn.Model(inp, out); f.compile("Adam", "binary_crossentropy", ["accuracy"], device="cpu")
    s = nn.Sequential([L.Conv2D(4, (3, 3), activation="relu", padding="same", input_shape=(12, 12, 2)), L.MaxPooling2D((2, 2)),
                       L.Flatten(), L.Dense(5, activation="relu"), L.Dense(1, activation="sigmoid")])
    s.compile("Adam", "binary_crossentropy", ["accuracy"], device="cpu")
    s.set_weights(f.get_weights())
    np.testing.assert_allclose(f.predict(x), s.predict(x), rtol=1e-6)
