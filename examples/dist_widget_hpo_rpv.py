#!/usr/bin/env python
"""Widget-monitored RPV HPO -- DistWidgetHPO_rpv.ipynb: 4 trials, 2 epochs (:113,125); the parameter dict holds 2-D arrays
`conv_sizes` / `fc_sizes` (one row per trial, :206-212) next to scalars, which the table shows stringified (the reference's
isinstance check was reversed, hpo_widgets.py:184)."""
import functools
import os
import time

import _common
import numpy as np


def build_and_train(path, n_train, n_valid, n_epochs, batch_size, device, conv_sizes, fc_sizes, lr, dropout, optimizer):
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    (xt, yt, _), (xv, yv, _), _ = rpv.load_dataset(path, n_train=n_train, n_valid=n_valid, n_test=0)
    model = rpv.build_model(xt.shape[1:], conv_sizes=conv_sizes, fc_sizes=fc_sizes, dropout=dropout, optimizer=optimizer,
                            lr=lr, device=device)
    history = rpv.train_model(model, xt, yt, xv, yv, batch_size=batch_size, n_epochs=n_epochs, verbose=2,
                              callbacks=[IPyParallelLogger()])
    return history.history


def main(argv=None):
    ap = _common.parser(__doc__, cluster=True, engines=None)
    ap.add_argument("--n-trials", type=int, default=4)
    ap.add_argument("--n-epochs", type=int, default=2)
    a = ap.parse_args(argv)
    n_train, n_valid, batch = 64000, 32000, 64
    if a.quick:
        a.n_trials, n_train, n_valid, batch = 2, 128, 64, 32
    from cori_intml_examples_b200.monitor import ParamSpanWidget, ModelPlot
    dev = _common.resolve_device(a.device)
    np.random.seed(0)
    n = a.n_trials
    conv = np.stack([np.random.choice([4, 8] if a.quick else [4, 8, 16, 32, 64], size=n),
                     np.random.choice([4, 8] if a.quick else [4, 8, 16, 32, 64], size=n),
                     np.random.choice([8, 16] if a.quick else [8, 16, 32, 64, 128], size=n)], axis=1)
    params = dict(conv_sizes=conv, fc_sizes=np.random.choice([16] if a.quick else [32, 64, 128, 256], size=(n, 1)),
                  lr=np.random.choice([0.0001, 0.001, 0.01], size=n), dropout=np.random.rand(n),
                  optimizer=np.random.choice(['Adadelta', 'Adam', 'Nadam'], size=n))
    with _common.cluster(a, default_engines=2) as c:
        compute = functools.partial(build_and_train, a.out_dir, n_train, n_valid, a.n_epochs, batch, dev)
        vis = functools.partial(ModelPlot, y=['loss', 'val_loss', 'acc', 'val_acc'], x='epoch', xlim=(0, a.n_epochs - 1), ylim=(0, 1))
        psw = ParamSpanWidget(compute, vis, params, client=c, interval=0.2 if a.quick else 1.0)
        psw.submit_computations()
        psw.wait()
        psw.select([0])
        print(psw.render_text())
        frame = psw.to_frame()
        psw.close()
        return frame


if __name__ == "__main__":
    main()
