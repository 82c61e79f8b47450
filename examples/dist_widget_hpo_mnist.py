#!/usr/bin/env python
"""Widget-monitored HPO -- DistWidgetHPO_mnist.ipynb: 8 trials (:146), `build_and_train` publishing epoch-wise history
through IPyParallelLogger, task function bound with functools.partial (:225-227), a ModelPlot factory with
y=['loss','val_loss','acc','val_acc'] (:229-235), ParamSpanWidget(compute_func, vis_func, params, ipp_cluster_id) +
submit_computations (:244-252).  There is no browser here: the table and the selected trial's plot are rendered as text (and
as an HTML file); "Stop selected" / "Restart selected" really abort / resubmit the tasks -- the reference's handlers were
stubs (hpo_widgets.py:352-364,386-391)."""
import functools
import os
import time

import _common
import numpy as np


def build_and_train(n_train, n_epochs, batch_size, device, h1, h2, h3, dropout, optimizer):
    from cori_intml_examples_b200.models import mnist
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    x_train, y_train, _, _ = mnist.load_data(n_train, 16)
    model = mnist.build_model(h1=h1, h2=h2, h3=h3, dropout=dropout, optimizer=optimizer, device=device)
    history = model.fit(x_train, y_train, batch_size=batch_size, epochs=n_epochs, validation_split=0.17,
                        callbacks=[IPyParallelLogger()], verbose=2)
    return history.history


def main(argv=None):
    ap = _common.parser(__doc__, cluster=True, engines=None)
    ap.add_argument("--n-trials", type=int, default=8)
    ap.add_argument("--n-epochs", type=int, default=16)
    a = ap.parse_args(argv)
    n_train, batch = 60000, 128
    if a.quick:
        a.n_trials, a.n_epochs, n_train, batch = 3, 3, 256, 64
    from cori_intml_examples_b200.monitor import ParamSpanWidget, ModelPlot
    dev = _common.resolve_device(a.device)
    np.random.seed(0)
    small = [4, 8] if a.quick else [4, 8, 16, 32, 64]
    params = dict(h1=np.random.choice(small, size=a.n_trials), h2=np.random.choice(small, size=a.n_trials),
                  h3=np.random.choice([8, 16] if a.quick else [8, 16, 32, 64, 128], size=a.n_trials),
                  dropout=np.random.rand(a.n_trials), optimizer=np.random.choice(['Adadelta', 'Adam', 'Nadam'], size=a.n_trials))
    with _common.cluster(a, default_engines=2) as c:
        compute = functools.partial(build_and_train, n_train, a.n_epochs, batch, dev)
        vis = functools.partial(ModelPlot, y=['loss', 'val_loss', 'acc', 'val_acc'], x='epoch', xlim=(0, a.n_epochs - 1), ylim=(0, 1),
                                xlabel='Epoch', title='Training history')
        psw = ParamSpanWidget(compute, vis, params, client=c, interval=0.2 if a.quick else 1.0)
        psw.submit_computations()
        if a.quick:
            # exercise the buttons: stop trial 0 as soon as it runs, then restart it
            t0 = time.time()
            while psw.rows[0]["status"] in ("Not started", "Submitted") and time.time() - t0 < 60:
                time.sleep(0.05)
            psw.select([0]); psw.stop_selected_models(); time.sleep(0.5); psw.restart_selected_models()
        while not psw.all_done():
            time.sleep(0.5 if a.quick else 5)
            if not a.quick:
                print(psw.render_text())
        psw.poll_once()
        print(psw.render_text())
        psw.select([a.n_trials - 1])
        print(psw.display_visualization().render_text())
        html = os.path.join(a.out_dir, "hpo_dashboard_%d.html" % os.getpid())
        with open(html, "w") as f:
            f.write(psw._repr_html_())
        print("dashboard written to", html)
        frame = psw.to_frame()
        psw.close()
        return frame


if __name__ == "__main__":
    main()
