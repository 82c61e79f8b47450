#!/usr/bin/env python
"""Synchronous data-parallel MNIST training driven from the client with %%px cells -- DistTrain_mnist.ipynb.

Cells reproduced: connect + ids (:44-62), `%%px` imports and hvd.init (:84-140), hyper-parameters batch 128 / 8 epochs
(:166-167), every rank loads the full data (:224), the inline Sequential 32/64/128 model with
Adadelta(1.0 * hvd.size()) wrapped in DistributedOptimizer (:294-310), fit with BroadcastGlobalVariablesCallback(0)
(:494-501), and the evaluation that prints the same test loss on every rank (:526-548)."""
import _common

SETUP = """
from cori_intml_examples_b200.parallel import dist as hvd
from cori_intml_examples_b200 import nn
from cori_intml_examples_b200.nn import layers, optimizers
from cori_intml_examples_b200.models import mnist
hvd.init()
print('MPI rank %i, local rank %i, size %i' % (hvd.rank(), hvd.local_rank(), hvd.size()))
"""

MODEL = """
batch_size, epochs = {batch}, {epochs}
x_train, y_train, x_test, y_test = mnist.load_data(n_train={n_train}, n_test={n_test})
model = nn.Sequential()
model.add(layers.Conv2D({h1}, kernel_size=(3, 3), activation='relu', input_shape=(28, 28, 1)))
model.add(layers.Conv2D({h2}, (3, 3), activation='relu'))
model.add(layers.MaxPooling2D(pool_size=(2, 2)))
model.add(layers.Dropout(0.25))
model.add(layers.Flatten())
model.add(layers.Dense({h3}, activation='relu'))
model.add(layers.Dropout(0.5))
model.add(layers.Dense(10, activation='softmax'))
opt = hvd.DistributedOptimizer(optimizers.Adadelta(1.0 * hvd.size()))
model.compile(loss='categorical_crossentropy', optimizer=opt, metrics=['accuracy'], device={device!r})
if hvd.rank() == 0:
    model.summary()
"""

TRAIN = """
callbacks = [hvd.callbacks.BroadcastGlobalVariablesCallback(0)]
history = model.fit(x_train, y_train, batch_size=batch_size, epochs=epochs, callbacks=callbacks,
                    verbose=2 if hvd.rank() == 0 else 0, validation_data=(x_test, y_test))
score = model.evaluate(x_test, y_test, verbose=0)
print('Test loss:', repr(score[0]))
print('Test accuracy:', score[1])
"""


def main(argv=None):
    ap = _common.parser(__doc__, cluster=True, engines=None)
    a = ap.parse_args(argv)
    dev = _common.resolve_device(a.device)
    cfg = dict(batch=128, epochs=8, n_train=60000, n_test=10000, h1=32, h2=64, h3=128, device=dev)
    if a.quick:
        cfg.update(batch=32, epochs=2, n_train=128, n_test=64, h1=4, h2=4, h3=8)
    import time
    with _common.cluster(a, shared_gpus=True, default_engines=2) as c, _common.Recorder(a, "dist_train_mnist") as rec:
        print("Worker IDs:", c.ids)
        dv = c[:]
        dv.setup_dist()
        dv.execute(SETUP)
        dv.execute(MODEL.format(**cfg))
        t0 = time.time()
        with rec.under_load():
            dv.execute(TRAIN)
        t_train = time.time() - t0
        scores = dv.get("score")
        rec["config"] = cfg; rec["ranks"] = len(c.ids); rec["train_wall_s"] = t_train
        rec["train_samples_per_s_all_ranks"] = len(c.ids) * cfg["n_train"] * cfg["epochs"] / t_train
        rec["test_loss_acc_per_rank"] = scores
        rec["history_rank0"] = c[0].get("history.history")
        print("test loss per rank:", [repr(s[0]) for s in scores])
        assert all(s[0] == scores[0][0] for s in scores), "replicas diverged"
        dv.execute("hvd.shutdown()", silent=True)
        return scores


if __name__ == "__main__":
    main()
