"""Every notebook workload of the reference exists as a script under examples/ and runs end to end on the CPU
tier in --quick mode (tiny synthetic data, process engines, gloo)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [
    ("train_rpv_single.py", [], ["Test loss:", "AUC:"]),                                   # Train_rpv.ipynb
    ("train_rpv_single.py", ["--legacy"], ["Test loss:", "Total params"]),                 #   ... its strided legacy model
    ("dist_train_mnist.py", ["--engines", "2"], ["Worker IDs: [0, 1]", "test loss per rank"]),      # DistTrain_mnist.ipynb
    ("dist_train_rpv.py", ["--engines", "2"], ["[stdout:1] MPI rank", "AUC:"]),              # DistTrain_rpv.ipynb
    ("hpo_mnist.py", [], ["Best validation accuracy", "Test accuracy"]),                   # HPO_mnist.ipynb
    ("dist_hpo_mnist.py", ["--engines", "2"], ["Tasks completed: 4 / 4", "Test accuracy"]),   # DistHPO_mnist.ipynb
    ("dist_hpo_rpv.py", ["--engines", "2"], ["trials/hour", "AUC:"]),                        # DistHPO_rpv.ipynb
    ("dist_widget_hpo_mnist.py", ["--engines", "2"], ["Ended Training", "dashboard written"]),  # DistWidgetHPO_mnist.ipynb
    ("dist_widget_hpo_rpv.py", ["--engines", "2"], ["Ended Training", "conv_sizes"]),        # DistWidgetHPO_rpv.ipynb
    ("gridsearchcv_mnist.py", [], ["baseline test", "Best:"]),                             # GridSearchCV_mnist.ipynb
    ("cray_hpo_mnist.py", ["--nodes", "2"], ["Best FoM"]),                                 # CrayHPO_mnist.ipynb
    ("cray_hpo_rpv.py", ["--nodes", "2", "--nodes-per-eval", "2"], ["Best FoM", "generation"]),  # CrayHPO_rpv.ipynb
]


@pytest.mark.parametrize("script,extra,expect", CASES, ids=[c[0][:-3] + ("-" + c[1][0].strip("-") if c[1] else "") for c in CASES])
def test_example_runs_quick(script, extra, expect, tmp_path):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2", CIB_CLUSTER_DIR=str(tmp_path / "clusters"), SCRATCH=str(tmp_path))
    cmd = [sys.executable, os.path.join(ROOT, "examples", script), "--quick", "--device", "cpu", "--out-dir", str(tmp_path)] + extra
    pr = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    assert pr.returncode == 0, (pr.stdout[-1500:] + "\n" + pr.stderr[-2500:])
    for s in expect:
        assert s in pr.stdout, "missing %r in output:\n%s" % (s, pr.stdout[-3000:])
