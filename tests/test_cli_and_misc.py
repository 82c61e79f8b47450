"""CLI trainers (FoM protocol), flat-name shims, mlextras, planner."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=ROOT)
    return subprocess.run([sys.executable] + args, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)


def test_train_rpv_cli_prints_fom():
    pr = _run(["train_rpv.py", "--n-train", "64", "--n-valid", "32", "--n-test", "16", "--n-epochs", "2", "--fom", "best", "--device", "cpu",
               "--h1", "4", "--h2", "4", "--h3", "8", "--h4", "8", "--batch-size", "32", "--lr-scaling", "linear"])
    assert pr.returncode == 0, pr.stderr[-1500:]
    out = pr.stdout
    assert "RPV classifier training: rank 0" in out and "Total params:" in out and "Test accuracy:" in out
    fom = [l for l in out.splitlines() if l.startswith("FoM: ")]
    assert len(fom) == 1 and 0 < float(fom[0].split()[1]) < 5


def test_train_mnist_cli_prints_fom():
    pr = _run(["train.py", "--epochs", "1", "--n-train", "128", "--n-test", "16", "--h1", "4", "--h2", "4", "--h3", "8", "--device", "cpu"])
    assert pr.returncode == 0, pr.stderr[-1500:]
    assert any(l.startswith("FoM: ") for l in pr.stdout.splitlines())


def test_flat_name_shims_have_reference_surface():
    sys.path.insert(0, ROOT)
    import mnist, rpv, mlextras, hpo_widgets, ipcluster_magics       # noqa: E401
    assert mnist.n_classes == 10 and (mnist.img_rows, mnist.img_cols) == (28, 28)
    for name in ("load_data", "build_model"):
        assert callable(getattr(mnist, name))
    for name in ("load_file", "load_dataset", "build_model", "train_model"):
        assert callable(getattr(rpv, name))
    import inspect
    sig = inspect.signature(rpv.train_model)
    assert list(sig.parameters)[:7] == ["model", "train_input", "train_labels", "valid_input", "valid_labels", "batch_size", "n_epochs"]
    assert sig.parameters["lr_reduce_patience"].default == 8 and sig.parameters["verbose"].default == 2
    assert inspect.signature(mnist.build_model).parameters["optimizer"].default == "Adadelta"
    assert hasattr(mlextras, "IPyParallelLogger") and hasattr(mlextras, "configure_session")
    for name in ("ModelPlot", "ParamSpanWidget", "ModelController", "ModelTaskData", "ModelPlotTable"):
        assert hasattr(hpo_widgets, name)
    assert callable(ipcluster_magics.ipcluster)


def test_ipyparallel_logger_publishes_reference_status_strings():
    from cori_intml_examples_b200.mlextras import IPyParallelLogger, configure_session
    from cori_intml_examples_b200.cluster import datapub
    from cori_intml_examples_b200.models import mnist
    x, y, _, _ = mnist.load_data(64, 8)
    m = mnist.build_model(4, 4, 8, 0.1, "Adam", device="cpu")
    seen = []
    orig = datapub.publish_data
    import cori_intml_examples_b200.mlextras as mx
    mx.publish_data = lambda d: seen.append((d["status"], d.get("epoch"), {k: list(v) for k, v in d["history"].items()}))
    try:
        m.fit(x, y, batch_size=32, epochs=2, validation_split=0.25, verbose=0, callbacks=[IPyParallelLogger()])
    finally:
        mx.publish_data = orig
    assert [s[0] for s in seen] == ["Begin Training", "Begin Epoch", "Ended Epoch", "Begin Epoch", "Ended Epoch", "Ended Training"]
    assert seen[-1][2]["epoch"] == [0, 1] and len(seen[-1][2]["val_acc"]) == 2
    os.environ["NUM_INTRA_THREADS"] = "3"
    cfg = configure_session()
    assert cfg.intra_op_parallelism_threads == 3 and cfg.inter_op_parallelism_threads == 2


def test_planner_padding_paths_and_pack_roundtrip():
    from cori_intml_examples_b200.models import mnist, rpv
    from cori_intml_examples_b200.ops.plan import build_plan, PlanError
    m = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=0.2, device="cpu")
    p = build_plan(m.layers, m.input_shape)
    kinds = [(b.kind, b.path) for b in p.blocks]
    assert kinds == [(0, 0), (0, 1), (0, 1), (1, 1), (2, 0)]            # conv1 CUDA-core, conv2/3 + dense tcgen05, head
    assert [b.Co for b in p.blocks[:3]] == [16, 32, 64] and p.blocks[2].drop_rate == pytest.approx(0.2)
    assert p.blocks[3].K == 8 * 8 * 64 and p.blocks[3].N == 128 and p.nparam % 512 == 0
    # arbitrary channel counts from genetic search (CrayHPO_rpv.ipynb: h1=13, h2=34, h4=153) are padded, not rejected
    m2 = rpv.build_model((64, 64, 1), [13, 34, 70], [153], device="cpu")
    p2 = build_plan(m2.layers, m2.input_shape)
    assert [b.Co for b in p2.blocks[:3]] == [16, 64, 96] and p2.blocks[3].N == 160
    w = m2.get_weights()
    flat = p2.pack(w)
    for a, b in zip(w, p2.unpack(flat)):
        np.testing.assert_array_equal(a, b)
    assert np.count_nonzero(flat) <= sum(x.size for x in w)            # padding is zero
    m3 = mnist.build_model(4, 8, 32, device="cpu")
    p3 = build_plan(m3.layers, m3.input_shape)
    assert p3.blocks[0].pool == 0 and p3.blocks[1].pool == 1 and p3.blocks[1].pad == 0 and p3.blocks[-1].head_kind == 1
    forced = build_plan(m.layers, m.input_shape, force_path=0)
    assert all(b.path == 0 for b in forced.blocks)
    from cori_intml_examples_b200 import nn
    from cori_intml_examples_b200.nn import layers as L
    odd = nn.Sequential([L.Flatten(input_shape=(8, 8, 1)), L.Dense(1, activation="sigmoid")])
    with pytest.raises(PlanError):
        build_plan(odd.layers, odd.input_shape)
