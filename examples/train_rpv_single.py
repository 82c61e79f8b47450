#!/usr/bin/env python
"""Single-process RPV training on one GPU -- the workload of Train_rpv.ipynb (cells :189-191 data, :204-218/:289-291 the
legacy strided model with 34,515,201 parameters, :248-249 batch 128 / 4 epochs, :304-312 fit log, then evaluation, metrics
and ROC)."""
import time

import _common
import numpy as np


def main(argv=None):
    ap = _common.parser(__doc__)
    ap.add_argument("--legacy", action="store_true", help="the 34.5M-parameter strided model of the old notebook")
    ap.add_argument("--n-train", type=int, default=64000)
    ap.add_argument("--n-valid", type=int, default=32000)
    ap.add_argument("--n-test", type=int, default=32000)
    ap.add_argument("--batch-size", type=int, default=128)
    ap.add_argument("--n-epochs", type=int, default=4)
    a = ap.parse_args(argv)
    if a.quick:
        a.n_train, a.n_valid, a.n_test, a.n_epochs, a.batch_size = 256, 64, 64, 2, 32
    from cori_intml_examples_b200.models import rpv
    from cori_intml_examples_b200.monitor import analysis
    dev = _common.resolve_device(a.device)
    (xt, yt, wt), (xv, yv, wv), (xs, ys, ws) = rpv.load_dataset(a.out_dir, a.n_train, a.n_valid, a.n_test)
    print("train", xt.shape, "valid", xv.shape, "test", xs.shape)
    if a.legacy:
        sizes = dict(h1=8, h2=8, h3=16, h4=16, h5=32) if a.quick else {}
        model = rpv.build_legacy_model(xt.shape[1:], optimizer="Adam", device=dev, **sizes)
    else:
        model = rpv.build_model(xt.shape[1:], conv_sizes=[16, 32, 64], fc_sizes=[128], dropout=0.2, optimizer="Adam", device=dev)
    model.summary()
    t0 = time.time()
    hist = rpv.train_model(model, xt, yt, xv, yv, batch_size=a.batch_size, n_epochs=a.n_epochs, verbose=2)
    print("training took %.1f s (%.1f s/epoch)" % (time.time() - t0, (time.time() - t0) / a.n_epochs))
    loss, acc = model.evaluate(xs, ys, verbose=0)
    print("Test loss: %.4f  Test accuracy: %.4f" % (loss, acc))
    out = model.predict(xs).squeeze(-1)
    print("unweighted:", analysis.summarize_metrics(ys, out))
    print("weighted:  ", analysis.summarize_metrics(ys, out, weights=ws))
    fpr, tpr, _ = analysis.roc_curve(ys, out)
    print("AUC: %.4f" % analysis.auc(fpr, tpr))
    return hist


if __name__ == "__main__":
    main()
