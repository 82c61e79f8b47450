#!/usr/bin/env python
"""Grid search with k-fold cross-validation -- GridSearchCV_mnist.ipynb: a baseline fit of the default model (37,562
parameters, 16 epochs, :194-250), then KerasClassifier(build_fn) inside a 3-fold grid over h1 in {8,16,32}, h2 in {16,32},
h3 in {16,32}, dropout in {0,.25,.5} = 36 candidates (:274-298), and the cv_results_ table (:363-386).  Works with
scikit-learn's own GridSearchCV (`--sklearn`) or with ours, which can also spread the 108 fits over a cluster view."""
import _common
import numpy as np


def main(argv=None):
    ap = _common.parser(__doc__, cluster=True, engines=0)
    ap.add_argument("--sklearn", action="store_true", help="use sklearn.model_selection.GridSearchCV (serial)")
    ap.add_argument("--n-epochs", type=int, default=16)
    a = ap.parse_args(argv)
    n_train, n_test, batch = 60000, 10000, 128
    grid = dict(h1=[8, 16, 32], h2=[16, 32], h3=[16, 32], dropout=[0, 0.25, 0.5])
    if a.quick:
        a.n_epochs, n_train, n_test, batch = 1, 240, 60, 40
        grid = dict(h1=[4, 8], dropout=[0, 0.5])
    from cori_intml_examples_b200 import hpo
    from cori_intml_examples_b200.models import mnist
    dev = _common.resolve_device(a.device)
    x_train, y_train, x_test, y_test = mnist.load_data(n_train, n_test)

    model = mnist.build_model(device=dev)
    model.summary()
    model.fit(x_train, y_train, batch_size=batch, epochs=a.n_epochs, verbose=2)
    print("baseline test [loss, acc]:", model.evaluate(x_test, y_test, verbose=0))
    print("predicted classes:", model.predict_classes(x_test[:10]).tolist())

    import functools
    clf = hpo.KerasClassifier(functools.partial(mnist.build_model, device=dev), batch_size=batch, epochs=a.n_epochs, verbose=0)
    if a.sklearn:
        from sklearn.model_selection import GridSearchCV
        gs = GridSearchCV(clf, grid, cv=3, n_jobs=1, verbose=2, return_train_score=True)
        gs.fit(x_train, y_train.argmax(1))
    elif a.engines:
        with _common.cluster(a) as c:
            gs = hpo.GridSearchCV(clf, grid, cv=3, verbose=2, view=c.load_balanced_view())
            gs.fit(x_train, y_train)
    else:
        gs = hpo.GridSearchCV(clf, grid, cv=3, verbose=2)
        gs.fit(x_train, y_train)
    print("Best: %f using %s" % (gs.best_score_, gs.best_params_))
    res = gs.cv_results_
    for p, m, t in zip(res["params"], res["mean_test_score"], res["mean_train_score"]):
        print("  test %.4f  train %.4f  %s" % (m, t, p))
    return gs


if __name__ == "__main__":
    main()
