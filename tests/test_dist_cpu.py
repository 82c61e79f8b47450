"""Synchronous data parallelism on the CPU tier: 2 processes, gloo, same semantics as the Horovod calls of the reference."""
import json
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import json, sys, numpy as np
    sys.path.insert(0, %r)
    from cori_intml_examples_b200.parallel import dist as hvd
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.utils import synthetic
    hvd.init()
    x, y, _ = synthetic.rpv_like(96, seed=0)                   # not sharded: every rank sees the same samples (reference convention)
    m = rpv.build_model((64, 64, 1), [4, 4, 8], [8], dropout=0.0, optimizer='Adam', lr=0.001 * hvd.size(), use_horovod=True, device='cpu')
    m.seed = 100 + hvd.rank()                                  # different shuffles per rank, like Horovod ranks
    w_before = [float(np.abs(w).sum()) for w in m.get_weights()]
    h = rpv.train_model(m, x[:64], y[:64], x[64:], y[64:], batch_size=16, n_epochs=2, use_horovod=True, verbose=0)
    out = {'rank': hvd.rank(), 'size': hvd.size(), 'local_rank': hvd.local_rank(), 'w_before': w_before,
           'w_sum': [float(np.abs(w).sum()) for w in m.get_weights()], 'w0': m.get_weights()[0].ravel()[:5].tolist(),
           'val_loss': h.history['val_loss'], 'loss': h.history['loss'], 'lr': h.history['lr'],
           'test': m.evaluate(x[64:], y[64:], verbose=0)}
    import os
    open(os.path.join(os.environ['CIB_TEST_OUT'], 'rank%%d.json' %% hvd.rank()), 'w').write(json.dumps(out))
    hvd.shutdown()
""" % ROOT)


def test_two_rank_gloo_data_parallel_replicas_stay_identical(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = 29650 + os.getpid() % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script)]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2", CIB_TEST_OUT=str(tmp_path))
    pr = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert pr.returncode == 0, pr.stderr[-2000:]
    res = [json.loads((tmp_path / ("rank%d.json" % r)).read_text()) for r in (0, 1)]
    assert [r["rank"] for r in res] == [0, 1] and all(r["size"] == 2 for r in res)
    a, b = res
    # BroadcastGlobalVariables(0) + averaged gradients => bit-identical replicas (DistTrain_mnist.ipynb:526-548)
    assert a["w_sum"] == b["w_sum"] and a["w0"] == b["w0"] and a["test"] == b["test"]
    assert a["w_before"] != b["w_before"]                       # they started from different random inits
    # MetricAverage: identical val_loss on all ranks, while the training loss is also averaged in the logs
    assert a["val_loss"] == b["val_loss"] and a["loss"] == b["loss"] and a["lr"] == b["lr"]


def test_single_rank_facade_is_a_noop():
    from cori_intml_examples_b200.parallel import dist as hvd
    import numpy as np
    hvd.init()
    assert hvd.size() == 1 and hvd.rank() == 0 and hvd.local_rank() == 0
    g = [np.ones(3, np.float32)]
    assert hvd.DistributedOptimizer("Adam").allreduce_gradients(g) is g
    assert hvd.allreduce(np.array([1.0, 2.0])).tolist() == [1.0, 2.0]
    x, y = hvd.shard(np.arange(10), np.arange(10) * 2)            # one rank: the whole array
    assert x.tolist() == list(range(10)) and y[-1] == 18
    cb = hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=0, verbose=1)       # the reference always passes 0 (no-op)
    hvd.shutdown()
