#!/usr/bin/env python
"""Synchronous data-parallel RPV training from the client -- DistTrain_rpv.ipynb: `%%px` hvd.init (:136-141), each rank
loads the same 64000/32000/32000 events (:196-207), conv [16,32,64] + fc [128], Adam, lr = 0.001 * hvd.size(), dropout 0.2,
batch 128, 4 epochs (:265-276), rpv.train_model(use_horovod=True) (:294-300), histories pulled with
c[0].get('history.epoch') / c[:].get('history.history') (:310-311), predictions on every rank (:405-408) and the
accuracy / purity / efficiency / ROC analysis on the client (:420-514)."""
import _common
import numpy as np

SETUP = """
from cori_intml_examples_b200.parallel import dist as hvd
from cori_intml_examples_b200.models import rpv
hvd.init()
print('MPI rank %i, local rank %i' % (hvd.rank(), hvd.local_rank()))
"""

LOAD = """
train_data, valid_data, test_data = rpv.load_dataset({path!r}, n_train={n_train}, n_valid={n_valid}, n_test={n_test})
train_input, train_labels, train_weights = train_data
valid_input, valid_labels, valid_weights = valid_data
test_input, test_labels, test_weights = test_data
print('train shape:', train_input.shape, 'Mean label:', train_labels.mean())
"""

BUILD = """
conv_sizes, fc_sizes = {conv}, {fc}
optimizer, lr, dropout = 'Adam', 0.001 * hvd.size(), 0.2
batch_size, n_epochs = {batch}, {epochs}
model = rpv.build_model(train_input.shape[1:], conv_sizes=conv_sizes, fc_sizes=fc_sizes, dropout=dropout,
                        optimizer=optimizer, lr=lr, use_horovod=True, device={device!r})
if hvd.rank() == 0:
    model.summary()
"""

TRAIN = """
history = rpv.train_model(model, train_input=train_input, train_labels=train_labels, valid_input=valid_input,
                          valid_labels=valid_labels, batch_size=batch_size, n_epochs=n_epochs, use_horovod=True,
                          lr_warmup_epochs={warmup}, verbose=2 if hvd.rank() == 0 else 0)
"""

PREDICT = """
test_output = model.predict(test_input).squeeze(-1)
"""


def main(argv=None):
    ap = _common.parser(__doc__, cluster=True, engines=None)
    ap.add_argument("--warmup-epochs", type=int, default=0,
                    help="LearningRateWarmupCallback epochs (rpv.py:91-92; the notebook passes 0).  With lr = 0.001 x 8 ranks and no warm-up the "
                         "synthetic set collapses to loss ln 2 -- on the fp32 CPU oracle at global batch 1024 as well; 2 epochs of warm-up train")
    a = ap.parse_args(argv)
    dev = _common.resolve_device(a.device)
    cfg = dict(path=a.out_dir, n_train=64000, n_valid=32000, n_test=32000, conv=[16, 32, 64], fc=[128], batch=128, epochs=4, device=dev)
    if a.quick:
        cfg.update(n_train=128, n_valid=64, n_test=64, conv=[4, 4, 8], fc=[8], batch=32, epochs=2)
    import time
    from cori_intml_examples_b200.monitor import analysis
    with _common.cluster(a, shared_gpus=True, default_engines=2) as c, _common.Recorder(a, "dist_train_rpv") as rec:
        print("Worker IDs:", c.ids)
        dv = c[:]
        dv.setup_dist()
        for cell in (SETUP, LOAD.format(**cfg), BUILD.format(**cfg)):
            dv.execute(cell)
        t0 = time.time()
        with rec.under_load():
            dv.execute(TRAIN.format(warmup=a.warmup_epochs))
        t_train = time.time() - t0
        dv.execute(PREDICT)
        rec["config"] = {k: v for k, v in cfg.items() if k != "path"}
        rec["ranks"] = len(c.ids); rec["lr_warmup_epochs"] = a.warmup_epochs
        rec["train_wall_s"] = t_train        # rpv.train_model: n_epochs x (training pass + validation pass), incl. engine creation + graph capture
        rec["train_samples_per_s_all_ranks"] = len(c.ids) * cfg["n_train"] * cfg["epochs"] / t_train
        epochs = c[0].get("history.epoch")
        histories = dv.get("history.history")
        for k in ("loss", "val_loss", "acc", "val_acc"):
            print(k, "per rank, last epoch:", ["%.4f" % h[k][-1] for h in histories])
        assert all(h["val_loss"] == histories[0]["val_loss"] for h in histories), "MetricAverage must give identical val_loss"
        labels, weights, outputs = c[0].get("test_labels"), c[0].get("test_weights"), c[0].get("test_output")
        print("epochs:", epochs)
        print("unweighted:", analysis.summarize_metrics(labels, outputs))
        print("weighted:  ", analysis.summarize_metrics(labels, outputs, weights=weights))
        fpr, tpr, _ = analysis.roc_curve(labels, outputs)
        print("AUC: %.4f" % analysis.auc(fpr, tpr))
        rec["history_per_rank"] = histories
        rec["test_metrics_unweighted"] = analysis.summarize_metrics(labels, outputs)
        rec["auc"] = float(analysis.auc(fpr, tpr))
        dv.execute("hvd.shutdown()", silent=True)
        return histories


if __name__ == "__main__":
    main()
