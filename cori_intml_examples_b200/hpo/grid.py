"""Grid search with k-fold cross-validation + a scikit-learn compatible estimator wrapper.

``KerasClassifier(build_fn, batch_size=, epochs=, verbose=, **build_params)`` has the surface of
``keras.wrappers.scikit_learn.KerasClassifier`` (reference GridSearchCV_mnist.ipynb:275-277): ``fit / predict /
predict_proba / score / get_params / set_params`` so that ``sklearn.model_selection.GridSearchCV`` can clone it.
``GridSearchCV`` here is a small built-in alternative producing the same ``cv_results_`` columns (``param_*``,
``mean_test_score``, ``mean_train_score``, ``rank_test_score``) that can also spread the (candidate, fold) fits over
a load-balanced cluster view instead of running serially (the reference ran ``n_jobs=1``, :308).
"""
import inspect
import itertools
import time

import numpy as np

try:
    from sklearn.base import BaseEstimator, ClassifierMixin
except Exception:       # pragma: no cover
    BaseEstimator = object

    class ClassifierMixin(object):
        pass


class KerasClassifier(BaseEstimator, ClassifierMixin):
    def __init__(self, build_fn=None, **sk_params):
        self.build_fn = build_fn
        self.sk_params = dict(sk_params)
        self.model = None

    _FIT_KEYS = ("batch_size", "epochs", "verbose", "validation_split", "shuffle", "callbacks")

    # ---- sklearn parameter protocol (clone() calls get_params / constructor)
    def get_params(self, deep=True):
        res = dict(self.sk_params)
        res["build_fn"] = self.build_fn
        return res

    def set_params(self, **params):
        for k, v in params.items():
            if k == "build_fn":
                self.build_fn = v
            else:
                self.sk_params[k] = v
        return self

    def _build_kwargs(self):
        names = set(inspect.signature(self.build_fn).parameters)
        return {k: v for k, v in self.sk_params.items() if k in names}

    def _fit_kwargs(self):
        return {k: v for k, v in self.sk_params.items() if k in self._FIT_KEYS}

    def fit(self, x, y, **kwargs):
        y = np.asarray(y)
        if y.ndim == 1 or y.shape[1] == 1:
            self.classes_ = np.unique(y)
        else:
            self.classes_ = np.arange(y.shape[1])
        self.model = self.build_fn(**self._build_kwargs())
        fk = self._fit_kwargs(); fk.update(kwargs)
        self.history_ = self.model.fit(x, self._targets(y), **fk)
        return self.history_

    def _targets(self, y):
        """Integer class labels are one-hot encoded for categorical_crossentropy models (as the Keras wrapper does)."""
        y = np.asarray(y)
        if self.model.loss == "categorical_crossentropy" and (y.ndim == 1 or y.shape[1] == 1):
            from ..utils.np_utils import to_categorical
            return to_categorical(y.ravel(), self.model.output_shape[-1])
        return y

    def predict_proba(self, x, **kw):
        p = self.model.predict(x, batch_size=self.sk_params.get("batch_size", 32))
        if p.shape[1] == 1:
            p = np.hstack([1 - p, p])
        return p

    def predict(self, x, **kw):
        return self.classes_[self.predict_proba(x).argmax(axis=1)]

    def score(self, x, y, **kw):
        out = self.model.evaluate(x, self._targets(y), batch_size=self.sk_params.get("batch_size", 32), verbose=0)
        return float(out[1]) if isinstance(out, (list, tuple)) else float(out)


def _kfold_indices(n, k):
    sizes = np.full(k, n // k, dtype=int); sizes[:n % k] += 1
    idx = np.arange(n); out = []; cur = 0
    for s in sizes:
        test = idx[cur:cur + s]; train = np.concatenate([idx[:cur], idx[cur + s:]])
        out.append((train, test)); cur += s
    return out


def _fit_and_score(build_fn, params, fit_params, x, y, train, test, return_train_score):
    est = KerasClassifier(build_fn, **dict(params, **fit_params))
    t0 = time.time()
    est.fit(x[train], y[train])
    fit_time = time.time() - t0
    test_score = est.score(x[test], y[test])
    train_score = est.score(x[train], y[train]) if return_train_score else float("nan")
    try:
        est.model.backend.close()
    except Exception:
        pass
    return test_score, train_score, fit_time


class GridSearchCV(object):
    """Exhaustive search over ``param_grid`` with ``cv``-fold cross-validation (default 3, as in the reference run)."""

    def __init__(self, estimator, param_grid, cv=3, verbose=0, return_train_score=True, view=None, refit=True):
        self.estimator, self.param_grid, self.cv, self.verbose = estimator, dict(param_grid), int(cv), verbose
        self.return_train_score, self.view, self.refit = return_train_score, view, refit

    def candidates(self):
        keys = sorted(self.param_grid)
        return [dict(zip(keys, vals)) for vals in itertools.product(*[self.param_grid[k] for k in keys])]

    def fit(self, x, y):
        x, y = np.asarray(x), np.asarray(y)
        cands = self.candidates()
        folds = _kfold_indices(len(x), self.cv)
        build_fn = self.estimator.build_fn
        fit_params = {k: v for k, v in self.estimator.sk_params.items() if k in KerasClassifier._FIT_KEYS}
        base = {k: v for k, v in self.estimator.sk_params.items() if k not in KerasClassifier._FIT_KEYS}
        if self.verbose:
            print("Fitting %d folds for each of %d candidates, totalling %d fits" % (self.cv, len(cands), self.cv * len(cands)))
        jobs = []
        for ci, cand in enumerate(cands):
            for fi, (tr, te) in enumerate(folds):
                args = (build_fn, dict(base, **cand), fit_params, x, y, tr, te, self.return_train_score)
                if self.view is not None:
                    jobs.append((ci, fi, self.view.apply(_fit_and_score, *args)))
                else:
                    res = _fit_and_score(*args)
                    if self.verbose:
                        print("[CV] %s, score=%.4f, total=%.1fs" % (", ".join("%s=%s" % kv for kv in sorted(cand.items())), res[0], res[2]))
                    jobs.append((ci, fi, res))
        test = np.zeros((len(cands), self.cv)); train = np.zeros_like(test); times = np.zeros_like(test)
        for ci, fi, r in jobs:
            r = r.get() if hasattr(r, "get") else r
            test[ci, fi], train[ci, fi], times[ci, fi] = r
        res = {"params": cands, "mean_test_score": test.mean(1), "std_test_score": test.std(1), "mean_train_score": train.mean(1),
               "std_train_score": train.std(1), "mean_fit_time": times.mean(1)}
        for k in sorted(self.param_grid):
            res["param_" + k] = np.array([c[k] for c in cands], dtype=object)
        for fi in range(self.cv):
            res["split%d_test_score" % fi] = test[:, fi]
        order = np.argsort(-res["mean_test_score"], kind="stable")
        rank = np.empty(len(cands), dtype=int); rank[order] = np.arange(1, len(cands) + 1)
        res["rank_test_score"] = rank
        self.cv_results_ = res
        self.best_index_ = int(order[0])
        self.best_params_ = cands[self.best_index_]
        self.best_score_ = float(res["mean_test_score"][self.best_index_])
        if self.refit:
            self.best_estimator_ = KerasClassifier(build_fn, **dict(base, **dict(self.best_params_, **fit_params)))
            self.best_estimator_.fit(x, y)
        return self

    def score(self, x, y):
        return self.best_estimator_.score(x, y)

    def predict(self, x):
        return self.best_estimator_.predict(x)
