#!/usr/bin/env python
"""Serial random-search HPO in one process -- HPO_mnist.ipynb: 16 trials, h1,h2 in {4..64}, h3 in {8..128}, dropout U[0,1),
optimizer in {Adadelta, Adam, Nadam} (:115-122), batch 128 / 16 epochs / validation_split 0.17 (:125-127), loop over trials
(:250-272), best trial by val_acc and its test accuracy (:1998-1999)."""
import _common
import numpy as np


def main(argv=None):
    ap = _common.parser(__doc__)
    ap.add_argument("--n-trials", type=int, default=16)
    ap.add_argument("--n-epochs", type=int, default=16)
    a = ap.parse_args(argv)
    n_train, n_test, batch = 60000, 10000, 128
    if a.quick:
        a.n_trials, a.n_epochs, n_train, n_test, batch = 3, 2, 256, 64, 64
    from cori_intml_examples_b200 import hpo
    from cori_intml_examples_b200.models import mnist
    from cori_intml_examples_b200.mlextras import configure_session
    configure_session()
    dev = _common.resolve_device(a.device)
    x_train, y_train, x_test, y_test = mnist.load_data(n_train, n_test)
    trials = hpo.sample_mnist_space(a.n_trials, seed=0)
    histories, models = [], []
    for i, hp in enumerate(trials):
        print("Hyperparameter trial %i: %s" % (i, hp))
        if a.quick:
            hp = dict(hp, h1=min(hp["h1"], 8), h2=min(hp["h2"], 8), h3=min(hp["h3"], 16))
        model = mnist.build_model(device=dev, **hp)
        h = model.fit(x_train, y_train, batch_size=batch, epochs=a.n_epochs, validation_split=0.17, verbose=2)
        histories.append(h.history); models.append(model)
    val_acc = np.array([max(h["val_acc"]) for h in histories])
    best = int(val_acc.argmax())
    print("Best validation accuracy: %.4f (trial %i: %s)" % (val_acc[best], best, trials[best]))
    score = models[best].evaluate(x_test, y_test, verbose=0)
    print("Test loss: %.4f  Test accuracy: %.4f" % tuple(score))
    return trials, histories


if __name__ == "__main__":
    main()
