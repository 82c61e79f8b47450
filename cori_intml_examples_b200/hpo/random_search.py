"""Random-search HPO: the notebook idiom of the reference as reusable functions.

``sample_mnist_space`` / ``sample_rpv_space`` reproduce the hyper-parameter tables of DistHPO_mnist.ipynb:136-145 and
DistHPO_rpv.ipynb:90-102 bit for bit under ``np.random.seed(0)`` (NumPy's legacy RandomState stream is stable).
``RandomSearch`` submits one load-balanced task per trial (``lv.apply``), exposes progress, and does the model
selection / reporting steps the notebooks do by hand (best / worst / top-k, per-task time, DistHPO_rpv.ipynb:217,340).
"""
import collections

import numpy as np


def sample_mnist_space(n_trials, seed=0):
    """h1,h2 in {4,8,16,32,64}; h3 in {8,...,128}; dropout U[0,1); optimizer in {Adadelta, Adam, Nadam}."""
    np.random.seed(seed)
    h1 = np.random.choice([4, 8, 16, 32, 64], size=n_trials)
    h2 = np.random.choice([4, 8, 16, 32, 64], size=n_trials)
    h3 = np.random.choice([8, 16, 32, 64, 128], size=n_trials)
    dropout = np.random.rand(n_trials)
    optimizer = np.random.choice(['Adadelta', 'Adam', 'Nadam'], size=n_trials)
    return [dict(h1=int(a), h2=int(b), h3=int(c), dropout=float(d), optimizer=str(o))
            for a, b, c, d, o in zip(h1, h2, h3, dropout, optimizer)]


def sample_rpv_space(n_trials, seed=0):
    """conv h1,h2 in {4..64}, h3 in {8..128}; fc in {32,64,128,256}; lr in {1e-4,1e-3,1e-2}; dropout U[0,1); optimizer."""
    np.random.seed(seed)
    h1 = np.random.choice([4, 8, 16, 32, 64], size=n_trials)
    h2 = np.random.choice([4, 8, 16, 32, 64], size=n_trials)
    h3 = np.random.choice([8, 16, 32, 64, 128], size=n_trials)
    fc = np.random.choice([32, 64, 128, 256], size=n_trials)
    lr = np.random.choice([0.0001, 0.001, 0.01], size=n_trials)
    dropout = np.random.rand(n_trials)
    optimizer = np.random.choice(['Adadelta', 'Adam', 'Nadam'], size=n_trials)
    return [dict(conv_sizes=[int(a), int(b), int(c)], fc_sizes=[int(f)], lr=float(l), dropout=float(d), optimizer=str(o))
            for a, b, c, f, l, d, o in zip(h1, h2, h3, fc, lr, dropout, optimizer)]


class RandomSearch(object):
    """Farm ``func(**params)`` over a load-balanced view; ``func`` returns a Keras-style ``history.history`` dict."""

    def __init__(self, view, func, trials, checkpoint_pattern=None, retries=0):
        self.view, self.func, self.trials = view, func, list(trials)
        self.checkpoint_pattern = checkpoint_pattern
        self.results = []
        view.retries = retries

    def submit(self):
        self.results = []
        for i, hp in enumerate(self.trials):
            kw = dict(hp)
            if self.checkpoint_pattern:
                kw["checkpoint_file"] = self.checkpoint_pattern % i
            self.results.append(self.view.apply(self.func, **kw))
        return self

    def n_done(self):
        return int(sum(ar.ready() for ar in self.results))

    def wait(self, timeout=None):
        for ar in self.results:
            ar.wait(timeout)
        return self

    def histories(self):
        return [ar.get() for ar in self.results]

    def scores(self, key="val_acc", reduce="max"):
        out = []
        for ar in self.results:
            try:
                h = ar.get()
                vals = h[key]
                out.append(max(vals) if reduce == "max" else min(vals) if reduce == "min" else vals[-1])
            except Exception:
                out.append(float("nan"))
        return np.array(out, dtype=float)

    def best(self, key="val_acc", reduce="max", maximize=True):
        s = self.scores(key, reduce)
        i = int(np.nanargmax(s) if maximize else np.nanargmin(s))
        return i, self.trials[i], float(s[i])

    def worst(self, key="val_acc", reduce="max", maximize=True):
        s = self.scores(key, reduce)
        i = int(np.nanargmin(s) if maximize else np.nanargmax(s))
        return i, self.trials[i], float(s[i])

    def top(self, k=5, key="val_acc", reduce="max", maximize=True):
        s = self.scores(key, reduce)
        order = np.argsort(-s if maximize else s)
        return [(int(i), self.trials[int(i)], float(s[int(i)])) for i in order[:k]]

    def task_times(self):
        """Seconds per task, ``completed - started`` (DistHPO_rpv.ipynb:217)."""
        return [(ar.completed - ar.started).total_seconds() if ar.completed and ar.started else float("nan") for ar in self.results]

    def table(self):
        rows = []
        s = self.scores()
        for i, (hp, ar) in enumerate(zip(self.trials, self.results)):
            row = collections.OrderedDict(trial=i); row.update(hp); row["score"] = s[i]; row["engine"] = ar.engine_id
            rows.append(row)
        try:
            import pandas as pd
            return pd.DataFrame(rows)
        except ImportError:
            return rows
