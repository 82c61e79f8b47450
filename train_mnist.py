#!/usr/bin/env python
"""MNIST CLI trainer, symmetric to train_rpv.py.  The reference's CrayHPO_mnist.ipynb:75 evaluates
``python train.py --epochs N --h1 .. --h2 .. --h3 .. --dropout .. --optimizer ..`` but never shipped that file
(SURVEY 2.8); this is it.  Prints ``FoM: <val_loss>`` (last epoch, or best with --fom best)."""
import argparse
import sys


def main(argv=None):
    p = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    p.add_argument('--epochs', type=int, default=4)
    p.add_argument('--h1', type=int, default=4)
    p.add_argument('--h2', type=int, default=8)
    p.add_argument('--h3', type=int, default=32)
    p.add_argument('--dropout', type=float, default=0.5)
    p.add_argument('--optimizer', default='Adadelta')
    p.add_argument('--batch-size', type=int, default=128)
    p.add_argument('--n-train', type=int, default=60000)
    p.add_argument('--n-test', type=int, default=10000)
    p.add_argument('--valid-frac', type=float, default=0.17)
    p.add_argument('--fom', choices=['best', 'last'], default='last')
    p.add_argument('--device', default='auto')
    p.add_argument('--verbose', type=int, default=2)
    a = p.parse_args(argv)
    from cori_intml_examples_b200.models import mnist
    x_train, y_train, x_test, y_test = mnist.load_data(a.n_train, a.n_test)
    model = mnist.build_model(a.h1, a.h2, a.h3, a.dropout, a.optimizer, device=a.device)
    h = model.fit(x_train, y_train, batch_size=a.batch_size, epochs=a.epochs, validation_split=a.valid_frac, verbose=a.verbose)
    vl = h.history['val_loss']
    sys.stdout.write('FoM: %r\n' % float(min(vl) if a.fom == 'best' else vl[-1])); sys.stdout.flush()
    return 0


if __name__ == '__main__':
    sys.exit(main())
