#!/usr/bin/env python
"""Batch / HPO entry point: (data-parallel) training of the ATLAS RPV CNN, printing ``FoM: <float>`` for the optimisers.

Flag-compatible with the reference ``train_rpv.py`` (argparse block at :16-32): --input-dir --n-train --n-valid
--n-test --h1..--h4 --dropout --lr --lr-scaling{linear} --optimizer --batch-size --n-epochs --fom{best,last}.
Run one process per GPU, e.g.
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 train_rpv.py --n-epochs 4 --fom best
(a plain ``python train_rpv.py`` is a one-rank world).  Extra flags: --channels (synthetic data), --device, --verbose.
"""
import argparse
import os
import socket
import sys

# (flag, type, default, extra argparse keywords) -- the first 15 rows are the reference's interface
FLAGS = [
    ("--input-dir", str, "/global/cscratch1/sd/sfarrell/atlas-rpv-images", {}),
    ("--n-train", int, 64000, {}), ("--n-valid", int, 32000, {}), ("--n-test", int, 0, {}),
    ("--h1", int, 16, {}), ("--h2", int, 32, {}), ("--h3", int, 64, {}), ("--h4", int, 128, {}),
    ("--dropout", float, 0.2, {}), ("--lr", float, 0.001, {}), ("--lr-scaling", str, None, {"choices": ["linear"]}),
    ("--optimizer", str, "Adam", {}), ("--batch-size", int, 64, {}), ("--n-epochs", int, 4, {}),
    ("--fom", str, None, {"choices": ["best", "last"]}),
    ("--channels", int, None, {"help": "channels of the synthetic images when the HDF5 files are absent"}),
    ("--device", str, "auto", {}), ("--verbose", int, 2, {}),
]


def parse_args(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    for flag, typ, default, extra in FLAGS:
        ap.add_argument(flag, type=typ, default=default, **extra)
    return ap.parse_args(argv)


def figure_of_merit(val_losses, mode):
    """``best`` = lowest validation loss of the run, ``last`` = the final epoch's (reference train_rpv.py:76-79)."""
    return min(val_losses) if mode == "best" else val_losses[-1]


def main(argv=None):
    a = parse_args(argv)
    if a.channels:
        os.environ["RPV_CHANNELS"] = str(a.channels)
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.parallel import dist as hvd
    hvd.init()
    chief = hvd.rank() == 0
    print("RPV classifier training: rank %i (local %i) of %i on %s" % (hvd.rank(), hvd.local_rank(), hvd.size(), socket.gethostname()))

    splits = rpv.load_dataset(a.input_dir, a.n_train, a.n_valid, a.n_test)
    for tag, (xs, ys, _) in zip(("train", "valid", "test"), splits):
        if len(xs):
            print("%s shape: %s  mean label: %.4f" % (tag, xs.shape, float(ys.mean())))
    (xt, yt, _), (xv, yv, _), (xs, ys, _) = splits

    world_lr = a.lr * (hvd.size() if a.lr_scaling == "linear" else 1)
    model = rpv.build_model(xt.shape[1:], conv_sizes=[a.h1, a.h2, a.h3], fc_sizes=[a.h4], dropout=a.dropout,
                            optimizer=a.optimizer, lr=world_lr, use_horovod=True, device=a.device)
    if chief:
        model.summary()
    hist = rpv.train_model(model, xt, yt, xv, yv, batch_size=a.batch_size, n_epochs=a.n_epochs, use_horovod=True,
                           verbose=a.verbose if chief else 0)
    if a.fom:
        # one write per line: several ranks share the pipe and the optimisers parse "FoM: <float>" from it
        sys.stdout.write("FoM: %r\n" % float(figure_of_merit(hist.history["val_loss"], a.fom)))
        sys.stdout.flush()
    if chief and a.n_test > 0:
        loss, acc = model.evaluate(xs, ys, verbose=0)
        print("Test loss:", loss)
        print("Test accuracy:", acc)
    hvd.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
