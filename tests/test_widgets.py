"""Headless HPO dashboard: table + plots updated from published data; stop / restart really work."""
import time

import pytest

import numpy as np

from cori_intml_examples_b200.cluster import LocalCluster
from cori_intml_examples_b200.monitor import ModelPlot, ModelPlotTable, ModelTaskData, ParamSpanWidget, analysis


def _trial(h1, dropout, conv_sizes, n_epochs=3, pause=0.0):
    import time
    from cori_intml_examples_b200.models import mnist
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    from cori_intml_examples_b200.nn.callbacks import LambdaCallback
    x, y, _, _ = mnist.load_data(n_train=128, n_test=8)
    model = mnist.build_model(h1, 4, 8, dropout, "Adam", device="cpu")
    slow = LambdaCallback(on_epoch_end=lambda e, l: time.sleep(pause))
    h = model.fit(x, y, batch_size=64, epochs=n_epochs, validation_split=0.25, verbose=0, callbacks=[IPyParallelLogger(), slow])
    return h.history


def test_plot_table_and_task_data():
    t = ModelPlotTable(["epoch", "loss"])
    t.append_row({"epoch": 0, "loss": 1.0}); t.append_row({"epoch": 1, "loss": 0.5})
    t.append_column("acc", [0.1, 0.9])
    assert t.to_dict() == {"epoch": [0, 1], "loss": [1.0, 0.5], "acc": [0.1, 0.9]} and t.num_rows() == 2
    d = ModelTaskData(["epoch", "loss"], ["status"])
    d.append_plot_data_row({"epoch": 0, "loss": 2.0})
    assert d.num_data_rows() == 1 and d.get_plot_data()["loss"] == [2.0]
    p = ModelPlot(y=["loss", "val_loss"], x="epoch", xlim=[0, 1], ylim=[0, 1], xlabel="Epoch", ylabel="Loss", title="t")
    p.update({"epoch": [0, 1, 2], "loss": [3.0, 2.0, 1.0], "val_loss": [2.5, 2.2, 1.9]})
    assert p.xlim == [0, 2] and p.ylim == [0, 3.0]                      # scales grow with the data
    assert "<polyline" in p.to_svg() and "loss" in p.render_text()


def test_param_span_widget_live_updates_stop_and_restart():
    params = {"h1": np.array([4, 8, 4]), "dropout": np.array([0.1, 0.2, 0.3]), "conv_sizes": np.array([[4, 8], [8, 8], [4, 4]])}
    import functools
    with LocalCluster(2, mode="process", gpus=[]) as cl:
        c = cl.client()
        psw = ParamSpanWidget(compute_func=functools.partial(_trial, n_epochs=3, pause=0.4),
                              vis_func=functools.partial(ModelPlot, y=["loss", "val_loss"], x="epoch", xlabel="Epoch", ylabel="Loss"),
                              params=params, client=c, interval=0.1)
        assert psw.columns == ["status", "epoch", "h1", "dropout", "conv_sizes", "loss", "val_loss", "acc", "val_acc"]
        assert psw.rows[0]["conv_sizes"] == "[4, 8]"                     # list-valued parameters are stringified for display
        psw.submit_computations()
        psw.select(0)
        # trial 2 is queued behind the two running ones: stop it before it starts, then stop a running one
        time.sleep(1.2)
        psw.select([2]); psw.stop_selected_models()
        assert psw.get_models_status()[2] == "Stopped"
        psw.select(0)
        assert psw.wait(60)
        st = psw.get_models_status()
        assert st[0] == "Ended Training" and st[1] == "Ended Training" and st[2] == "Stopped"
        assert psw.model_data[0].num_data_rows() == 3                    # one row per epoch, no duplicates
        assert psw.rows[0]["epoch"] == 2 and isinstance(psw.rows[0]["val_acc"], float)
        assert "Ended Training" in psw.render_text() and "<table>" in psw._repr_html_()
        assert len(psw.model_plots[0].series["loss"][1]) == 3
        psw.select([2]); psw.restart_selected_models()
        assert psw.wait(60) and psw.get_models_status()[2] == "Ended Training"
        assert psw.to_frame().shape == (3, 9)
        psw.close()
        c.close()


def test_analysis_metrics_match_sklearn():
    from sklearn import metrics
    rng = np.random.RandomState(0)
    y = (rng.rand(500) > 0.5).astype(np.float32)
    out = np.clip(y * 0.6 + rng.rand(500) * 0.5, 0, 1)
    w = rng.rand(500) + 0.5
    m = analysis.summarize_metrics(y, out, weights=w)
    pred = out > 0.5
    assert abs(m["accuracy"] - metrics.accuracy_score(y, pred, sample_weight=w)) < 1e-9
    assert abs(m["purity"] - metrics.precision_score(y, pred, sample_weight=w)) < 1e-9
    assert abs(m["efficiency"] - metrics.recall_score(y, pred, sample_weight=w)) < 1e-9
    fpr, tpr, _ = analysis.roc_curve(y, out)
    assert abs(analysis.auc(fpr, tpr) - metrics.roc_auc_score(y, out)) < 1e-9
    h = analysis.output_histograms(y, out, bins=10)
    assert h["signal"].sum() == y.sum() and h["background"].sum() == 500 - y.sum()


def test_optional_ipywidgets_view_degrades_cleanly():
    import importlib.util
    from cori_intml_examples_b200.monitor import ParamSpanWidget, ModelPlot
    class _View(object):
        retries = 0

    class _Client(object):
        def load_balanced_view(self):
            return _View()
    psw = ParamSpanWidget(lambda **kw: None, lambda: ModelPlot(y=["loss"]), {"a": [1, 2]}, client=_Client(), autostart=False)
    if importlib.util.find_spec("ipywidgets") is None:
        with pytest.raises(ImportError, match="ipywidgets"):
            psw.ipywidget()
    else:
        assert psw.ipywidget() is not None
    psw.close()
