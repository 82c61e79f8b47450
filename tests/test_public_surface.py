"""the public surface a user of the reference relies on exists, under the reference's flat module names,
with the reference's signatures."""
import inspect


def _params(f):
    return list(inspect.signature(f).parameters)


def test_flat_modules_and_signatures():
    import hpo_widgets
    import ipcluster_magics
    import mlextras
    import mnist
    import rpv
    assert _params(mnist.build_model)[:5] == ["h1", "h2", "h3", "dropout", "optimizer"]
    d = inspect.signature(mnist.build_model).parameters
    assert (d["h1"].default, d["h2"].default, d["h3"].default, d["dropout"].default, d["optimizer"].default) == (4, 8, 32, 0.5, "Adadelta")
    assert mnist.n_classes == 10 and mnist.img_rows == mnist.img_cols == 28 and callable(mnist.load_data)
    assert _params(rpv.load_file) == ["filename", "n_samples"]
    d = inspect.signature(rpv.load_dataset).parameters
    assert (d["n_train"].default, d["n_valid"].default, d["n_test"].default) == (412416, 137471, 137471)
    assert _params(rpv.build_model)[:7] == ["input_shape", "conv_sizes", "fc_sizes", "dropout", "optimizer", "lr", "use_horovod"]
    assert _params(rpv.train_model)[:13] == ["model", "train_input", "train_labels", "valid_input", "valid_labels", "batch_size", "n_epochs",
                                             "lr_warmup_epochs", "lr_reduce_patience", "checkpoint_file", "use_horovod", "verbose", "callbacks"]
    assert callable(mlextras.IPyParallelLogger) and callable(mlextras.configure_session)
    for name in ("ModelPlot", "ParamSpanWidget", "ModelController", "ModelTaskData", "ModelPlotTable"):
        assert hasattr(hpo_widgets, name), name
    assert _params(hpo_widgets.ModelPlot.__init__)[1:8] == ["y", "x", "xlim", "ylim", "xlabel", "ylabel", "title"]
    assert _params(hpo_widgets.ParamSpanWidget.__init__)[1:8] == ["compute_func", "vis_func", "params", "columns", "ipp_cluster_id", "output_layout", "qgrid_layout"]
    for m in ("submit_computations", "stop_selected_models", "restart_selected_models", "get_models_status"):
        assert callable(getattr(hpo_widgets.ParamSpanWidget, m))
    for m in ("start_model", "stop_model", "restart_model", "set_model_completed", "get_completed_models", "get_running_models"):
        assert callable(getattr(hpo_widgets.ModelController, m))
    assert hasattr(ipcluster_magics, "IPClusterMagics") and "Usage" in ipcluster_magics.IPClusterMagics.__doc__


def test_model_dist_cluster_and_hpo_names():
    from cori_intml_examples_b200 import cluster, hpo, nn
    from cori_intml_examples_b200.parallel import dist as hvd
    for m in ("fit", "evaluate", "predict", "predict_classes", "summary", "save", "fit_stream", "get_weights", "set_weights"):
        assert callable(getattr(nn.Model, m)), m
    assert callable(nn.load_model)
    for f in ("init", "rank", "local_rank", "size", "DistributedOptimizer", "allreduce", "broadcast"):
        assert callable(getattr(hvd, f)), f
    for cb in ("BroadcastGlobalVariablesCallback", "MetricAverageCallback", "LearningRateWarmupCallback"):
        assert hasattr(hvd.callbacks, cb), cb
    for n in ("Client", "LocalCluster", "publish_data", "RemoteError", "EngineError", "TaskAborted"):
        assert hasattr(cluster, n), n
    from cori_intml_examples_b200.cluster.client import AsyncResult, DirectView, LoadBalancedView
    for a in ("ready", "done", "get", "abort", "stdout", "stderr", "started", "completed", "metadata", "data"):
        assert hasattr(AsyncResult, a), a
    assert callable(DirectView.execute) and callable(DirectView.get) and callable(DirectView.setup_dist) and callable(LoadBalancedView.apply)
    for n in ("Params", "Evaluator", "GeneticOptimizer", "RandomSearch", "GridSearchCV", "KerasClassifier", "sample_mnist_space", "sample_rpv_space"):
        assert hasattr(hpo, n), n
    assert _params(hpo.Evaluator.__init__)[1:7] == ["cmd", "nodes", "nodes_per_eval", "alloc_args", "launcher", "verbose"]
    assert _params(hpo.GeneticOptimizer.__init__)[1:9] == ["evaluator", "pop_size", "num_demes", "generations", "mutation_rate", "crossover_rate", "verbose", "log_fn"]


def test_train_rpv_cli_has_the_reference_flags():
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "train_rpv.py"), "--help"], capture_output=True, text=True, timeout=120).stdout
    for flag in ("--input-dir", "--n-train", "--n-valid", "--n-test", "--h1", "--h2", "--h3", "--h4", "--dropout", "--lr", "--lr-scaling",
                 "--optimizer", "--batch-size", "--n-epochs", "--fom"):
        assert flag in out, flag
