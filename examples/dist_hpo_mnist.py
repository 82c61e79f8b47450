#!/usr/bin/env python
"""Random-search HPO farmed over a load-balanced view -- DistHPO_mnist.ipynb: 32 trials under np.random.seed(0) (:136-145),
`build_and_train` with all imports inside and a ModelCheckpoint per trial (:170-191), `lv.apply` per trial with
`model_%i.h5` (:240-255), progress via `.ready()` (:272), captured stdout of one task (:333-335), histories with `.get()`
(:344), best/worst trial (:360-391), reload of the best checkpoint with load_model and test evaluation (:540-564)."""
import os
import time

import _common
import numpy as np


def build_and_train(n_train, n_test, n_epochs, batch_size, device, h1, h2, h3, dropout, optimizer, checkpoint_file=None,
                    valid_frac=0.17):
    """Runs on an engine, in an empty namespace: every import lives here."""
    from cori_intml_examples_b200.models import mnist
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    from cori_intml_examples_b200.nn.callbacks import ModelCheckpoint
    x_train, y_train, _, _ = mnist.load_data(n_train, n_test)
    model = mnist.build_model(h1=h1, h2=h2, h3=h3, dropout=dropout, optimizer=optimizer, device=device)
    callbacks = [IPyParallelLogger()]
    if checkpoint_file is not None:
        callbacks.append(ModelCheckpoint(checkpoint_file))
    history = model.fit(x_train, y_train, batch_size=batch_size, epochs=n_epochs, validation_split=valid_frac,
                        callbacks=callbacks, verbose=2)
    return history.history


def main(argv=None):
    import functools
    ap = _common.parser(__doc__, cluster=True, engines=None)
    ap.add_argument("--n-trials", type=int, default=32)
    ap.add_argument("--n-epochs", type=int, default=16)
    a = ap.parse_args(argv)
    n_train, n_test, batch = 60000, 10000, 128
    if a.quick:
        a.n_trials, a.n_epochs, n_train, n_test, batch = 4, 2, 256, 64, 64
    from cori_intml_examples_b200 import hpo, nn
    from cori_intml_examples_b200.models import mnist
    dev = _common.resolve_device(a.device)
    trials = hpo.sample_mnist_space(a.n_trials, seed=0)
    if a.quick:
        trials = [dict(t, h1=min(t["h1"], 8), h2=min(t["h2"], 8), h3=min(t["h3"], 16)) for t in trials]
    ckpt_dir = os.path.join(a.out_dir, "mnist-hpo-%d" % os.getpid())
    os.makedirs(ckpt_dir, exist_ok=True)
    with _common.cluster(a, default_engines=2) as c:
        print("Worker IDs:", c.ids)
        lv = c.load_balanced_view()
        func = functools.partial(build_and_train, n_train, n_test, a.n_epochs, batch, dev)
        t0 = time.time()
        rs = hpo.RandomSearch(lv, func, trials, checkpoint_pattern=os.path.join(ckpt_dir, "model_%i.h5")).submit()
        while rs.n_done() < len(trials):
            print("Tasks completed: %i / %i" % (rs.n_done(), len(trials)))
            time.sleep(0.5 if a.quick else 5)
        wall = time.time() - t0
        print("Tasks completed: %i / %i in %.1f s  (%.0f trials/hour)" % (rs.n_done(), len(trials), wall, 3600 * len(trials) / wall))
        print("--- stdout of trial 0 ---")
        print(rs.results[0].stdout[-600:])
        i, hp, score = rs.best()
        j, hpw, sw = rs.worst()
        print("Best  validation accuracy %.4f: trial %i %s" % (score, i, hp))
        print("Worst validation accuracy %.4f: trial %i %s" % (sw, j, hpw))
        print("task times [s]:", ["%.1f" % t for t in rs.task_times()])
        best_model = nn.load_model(os.path.join(ckpt_dir, "model_%i.h5" % i), device=dev)
        _, _, x_test, y_test = mnist.load_data(n_train, n_test)
        score = best_model.evaluate(x_test, y_test, verbose=0)
        print("Test loss: %.4f  Test accuracy: %.4f" % tuple(score))
        return rs.histories()


if __name__ == "__main__":
    main()
