#!/usr/bin/env python
"""Random-search HPO of the RPV classifier over the cluster -- DistHPO_rpv.ipynb: 32 trials, conv sizes / fc size / lr /
dropout / optimizer sampled under seed 0 (:90-102), batch 64 / 16 epochs (:105-106), each task loads the dataset itself
(:136-137), per-task times from `.started/.completed` (:217,235-237), top-5 table (:340-343), reload of the best model and
ROC / metrics on the test set (:399-)."""
import functools
import os
import time

import _common
import numpy as np


def build_and_train(path, n_train, n_valid, n_epochs, batch_size, device, conv_sizes, fc_sizes, lr, dropout, optimizer,
                    checkpoint_file=None):
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    (xt, yt, _), (xv, yv, _), _ = rpv.load_dataset(path, n_train=n_train, n_valid=n_valid, n_test=0)
    model = rpv.build_model(xt.shape[1:], conv_sizes=conv_sizes, fc_sizes=fc_sizes, dropout=dropout, optimizer=optimizer,
                            lr=lr, device=device)
    history = rpv.train_model(model, xt, yt, xv, yv, batch_size=batch_size, n_epochs=n_epochs, checkpoint_file=checkpoint_file,
                              verbose=2, callbacks=[IPyParallelLogger()])
    return history.history


def main(argv=None):
    ap = _common.parser(__doc__, cluster=True, engines=None)
    ap.add_argument("--n-trials", type=int, default=32)
    ap.add_argument("--n-epochs", type=int, default=16)
    a = ap.parse_args(argv)
    n_train, n_valid, n_test, batch = 64000, 32000, 32000, 64
    if a.quick:
        a.n_trials, a.n_epochs, n_train, n_valid, n_test, batch = 3, 2, 128, 64, 64, 32
    from cori_intml_examples_b200 import hpo, nn
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.monitor import analysis
    dev = _common.resolve_device(a.device)
    trials = hpo.sample_rpv_space(a.n_trials, seed=0)
    if a.quick:
        trials = [dict(t, conv_sizes=[min(v, 8) for v in t["conv_sizes"]], fc_sizes=[min(v, 16) for v in t["fc_sizes"]]) for t in trials]
    ckpt_dir = os.path.join(a.out_dir, "rpv-hpo-%d" % os.getpid())
    os.makedirs(ckpt_dir, exist_ok=True)
    with _common.cluster(a, default_engines=2) as c:
        lv = c.load_balanced_view()
        func = functools.partial(build_and_train, a.out_dir, n_train, n_valid, a.n_epochs, batch, dev)
        t0 = time.time()
        rs = hpo.RandomSearch(lv, func, trials, checkpoint_pattern=os.path.join(ckpt_dir, "model_%i.h5")).submit()
        rs.wait()
        wall = time.time() - t0
        print("%i trials in %.1f s  (%.0f trials/hour)" % (len(trials), wall, 3600 * len(trials) / wall))
        print("task times [s]: mean %.1f  max %.1f" % (np.nanmean(rs.task_times()), np.nanmax(rs.task_times())))
        print("Top trials by validation accuracy:")
        for i, hp, s in rs.top(5):
            print("  trial %2i  val_acc %.4f  %s" % (i, s, hp))
        i, hp, s = rs.best()
        model = nn.load_model(os.path.join(ckpt_dir, "model_%i.h5" % i), device=dev)
        _, _, (xs, ys, ws) = rpv.load_dataset(a.out_dir, n_train=0, n_valid=0, n_test=n_test)
        out = model.predict(xs).squeeze(-1)
        print("best trial %i on the test set:" % i, analysis.summarize_metrics(ys, out, weights=ws))
        fpr, tpr, _ = analysis.roc_curve(ys, out)
        print("AUC: %.4f" % analysis.auc(fpr, tpr))
        return rs.histories()


if __name__ == "__main__":
    main()
