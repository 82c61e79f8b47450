"""HPO drivers: random search (reference search points), grid search + CV, genetic search over a CLI."""
import os
import sys
import textwrap

import numpy as np
import pytest

from cori_intml_examples_b200 import hpo
from cori_intml_examples_b200.cluster import LocalCluster


def test_random_search_points_match_reference_table():
    t = hpo.sample_mnist_space(32, seed=0)
    # DistHPO_mnist.ipynb:203-234 printed trial table under np.random.seed(0)
    assert (t[0]["h1"], t[0]["h2"], t[0]["h3"], t[0]["optimizer"]) == (64, 8, 128, "Nadam") and abs(t[0]["dropout"] - 0.020) < 5e-4
    assert abs(t[2]["dropout"] - 0.980) < 5e-4                     # the degenerate trial (:205)
    assert (t[24]["h1"], t[24]["h2"], t[24]["h3"], t[24]["optimizer"]) == (16, 8, 128, "Adam") and abs(t[24]["dropout"] - 0.370) < 5e-4
    r = hpo.sample_rpv_space(4, seed=0)
    assert set(r[0]) == {"conv_sizes", "fc_sizes", "lr", "dropout", "optimizer"} and len(r[0]["conv_sizes"]) == 3


def _build_and_train(h1, h2, h3, dropout, optimizer, checkpoint_file=None):
    # imports inside the function: it runs in an empty engine namespace (DistHPO_mnist.ipynb:173-175)
    from cori_intml_examples_b200.models import mnist
    from cori_intml_examples_b200.mlextras import IPyParallelLogger
    from cori_intml_examples_b200.nn.callbacks import ModelCheckpoint
    x_train, y_train, _, _ = mnist.load_data(n_train=256, n_test=16)
    model = mnist.build_model(h1, h2, h3, dropout, optimizer, device="cpu")
    cbs = [IPyParallelLogger()] + ([ModelCheckpoint(checkpoint_file)] if checkpoint_file else [])
    h = model.fit(x_train, y_train, batch_size=64, epochs=2, validation_split=0.25, verbose=2, callbacks=cbs)
    return h.history


def test_random_search_over_cluster_with_checkpoints(tmp_path):
    trials = [dict(h1=4, h2=4, h3=8, dropout=0.1, optimizer="Adam"), dict(h1=4, h2=8, h3=8, dropout=0.5, optimizer="Nadam"),
              dict(h1=8, h2=4, h3=16, dropout=0.3, optimizer="Adadelta")]
    with LocalCluster(2, mode="process", gpus=[]) as cl:
        c = cl.client()
        rs = hpo.RandomSearch(c.load_balanced_view(), _build_and_train, trials, checkpoint_pattern=str(tmp_path / "model_%i.h5")).submit()
        rs.wait(120)
        assert rs.n_done() == 3
        hist = rs.histories()
        assert all(sorted(h) == ["acc", "loss", "val_acc", "val_loss"] and len(h["loss"]) == 2 for h in hist)
        i, hp, score = rs.best()
        assert hp == trials[i] and 0 <= score <= 1 and len(rs.top(2)) == 2
        assert all(t > 0 for t in rs.task_times())
        assert "Epoch 2/2" in rs.results[0].stdout                       # per-task captured output (DistHPO_mnist.ipynb:333)
        assert rs.results[0].data["status"] == "Ended Training" and rs.results[0].data["history"]["epoch"] == [0, 1]
        from cori_intml_examples_b200 import nn
        best = nn.load_model(str(tmp_path / ("model_%i.h5" % i)), device="cpu")     # DistHPO_mnist.ipynb:540-551
        from cori_intml_examples_b200.models import mnist
        _, _, xt, yt = mnist.load_data(16, 64)
        assert len(best.evaluate(xt, yt, verbose=0)) == 2
        c.close()


def test_grid_search_cv_and_sklearn_compat():
    from cori_intml_examples_b200.models import mnist
    x, y, _, _ = mnist.load_data(n_train=240, n_test=8)
    labels = y.argmax(1)

    def build(h1=4, h2=4, h3=8, dropout=0.1):
        return mnist.build_model(h1, h2, h3, dropout, "Adam", device="cpu")
    clf = hpo.KerasClassifier(build, batch_size=64, epochs=2, verbose=0)
    gs = hpo.GridSearchCV(clf, {"h1": [4, 8], "dropout": [0.0, 0.25]}, cv=3, refit=False)
    gs.fit(x, y)
    r = gs.cv_results_
    assert len(r["params"]) == 4 and set(r) >= {"param_h1", "param_dropout", "mean_test_score", "mean_train_score", "rank_test_score"}
    assert sorted(r["rank_test_score"]) == [1, 2, 3, 4] and gs.best_params_ in r["params"]
    # the estimator also works under scikit-learn's own GridSearchCV (GridSearchCV_mnist.ipynb:274-284)
    from sklearn.base import clone
    from sklearn.model_selection import GridSearchCV as SkGrid
    c2 = clone(clf)
    assert c2.get_params()["epochs"] == 2
    sk = SkGrid(hpo.KerasClassifier(build, batch_size=64, epochs=1, verbose=0), {"h1": [4, 8]}, cv=2, n_jobs=1, refit=False)
    sk.fit(x, labels)
    assert len(sk.cv_results_["mean_test_score"]) == 2


FAKE_TRAINER = textwrap.dedent("""
    import argparse
    p = argparse.ArgumentParser()
    p.add_argument('--h1', type=int); p.add_argument('--lr', type=float); p.add_argument('--optimizer'); p.add_argument('--n-epochs', type=int)
    a = p.parse_args()
    fom = (a.h1 - 37) ** 2 / 1000.0 + abs(a.lr - 0.01) * 10 + (0.0 if a.optimizer == 'Adam' else 0.5)
    print('some training output')
    print('FoM:', fom)
""")


def test_genetic_optimizer_cli_protocol_and_logs(tmp_path):
    script = tmp_path / "fake_train.py"
    script.write_text(FAKE_TRAINER)
    params = hpo.Params([["--h1", 16, (4, 64)], ["--lr", 0.001, (0.0001, 0.05)], ["--optimizer", "Adam", ["Adam", "Nadam", "Adadelta"]]])
    ev = hpo.Evaluator("python %s --n-epochs 4" % script, nodes=4, nodes_per_eval=1, verbose=False, env={"CIB_HPO_CPU": "1"})
    log = str(tmp_path / "hpo.log")
    opt = hpo.GeneticOptimizer(ev, pop_size=6, num_demes=2, generations=4, mutation_rate=0.2, crossover_rate=0.33, log_fn=log, seed=1)
    best = opt.optimize(params)
    assert set(best) == {"--h1", "--lr", "--optimizer"} and isinstance(best["--h1"], int)
    assert ev.n_evals == 6 * 2 * 4 and ev.n_concurrent == 4
    foms = [h["best_fom"] for h in opt.history]
    assert foms[-1] <= foms[0] and opt.best_fom == min(e["fom"] for e in opt.evaluations)
    first = opt.evaluations[0]
    assert first["params"] == [16, 0.001, "Adam"]                      # generation 0 contains the defaults
    head = open(log).readline().split()
    assert head == ["generation", "epoch", "best_fom", "avg_fom", "checkpoint_in", "checkpoint_out", "--h1", "--lr", "--optimizer"]
    assert len(open(log).read().strip().splitlines()) == 1 + 4
    d1 = open(str(tmp_path / "Deme1_hpo.log")).read().strip().splitlines()
    assert d1[0].split()[:4] == ["generation", "tag", "fitness", "FoM"] and len(d1) == 1 + 6 * 4
    import pandas as pd
    df = pd.read_csv(log, sep=r"\s+")                                   # CrayHPO_rpv.ipynb:342 analysis idiom
    assert list(df["generation"]) == [0, 1, 2, 3]


def test_genetic_over_real_cli_trainer(tmp_path):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    params = hpo.Params([["--h1", 4, (4, 8)], ["--dropout", 0.2, (0.0, 0.5)]])
    ev = hpo.Evaluator("%s %s --n-train 64 --n-valid 32 --n-epochs 1 --fom best --device cpu --h2 4 --h3 4 --h4 8 --batch-size 32"
                       % (sys.executable, os.path.join(root, "train_rpv.py")), nodes=2, nodes_per_eval=1, env={"CIB_HPO_CPU": "1", "PYTHONPATH": root})
    opt = hpo.GeneticOptimizer(ev, pop_size=2, num_demes=1, generations=1, seed=0)
    best = opt.optimize(params)
    assert best is not None and np.isfinite(opt.best_fom) and 0 < opt.best_fom < 2


def test_genetic_over_cluster_engines_with_data_parallel_groups():
    """Evaluator(launcher='cluster'): the evaluations run inside long-lived engines grouped into persistent data-parallel
    communicators (2 groups x 2 engines here, gloo on CPU) -- the hybrid task-parallel-over-DP strategy of CrayHPO_rpv.ipynb:145-151
    without a process start per evaluation.  Every engine of a group prints the same FoM (MetricAverageCallback)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cl = LocalCluster(n_engines=4, mode="process", gpus=[], env={"CIB_HPO_CPU": "1"}).start()
    try:
        c = cl.client()
        params = hpo.Params([["--h1", 4, (4, 8)], ["--dropout", 0.2, (0.0, 0.5)]])
        ev = hpo.Evaluator("python %s --n-train 64 --n-valid 32 --n-epochs 1 --fom best --device cpu --h2 4 --h3 4 --h4 8 --batch-size 32 --verbose 0"
                           % os.path.join(root, "train_rpv.py"), nodes=4, nodes_per_eval=2, launcher="cluster", client=c, timeout=300)
        opt = hpo.GeneticOptimizer(ev, pop_size=4, num_demes=1, generations=2, seed=0)
        best = opt.optimize(params)
        assert best is not None and np.isfinite(opt.best_fom) and 0 < opt.best_fom < 2
        assert ev.n_evals == 8 and ev.n_concurrent == 2 and len(ev._groups) == 2 and all(len(g) == 2 for g in ev._groups)
        assert all(np.isfinite(e["fom"]) for e in opt.evaluations)
        # the groups' communicators survived the evaluations (each training called hvd.shutdown(), which persistent groups ignore)
        sizes = c[:].apply_sync(lambda: __import__("cori_intml_examples_b200.parallel.dist", fromlist=["size"]).size())
        assert sizes == [2, 2, 2, 2]
    finally:
        cl.stop()
