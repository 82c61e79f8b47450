#!/usr/bin/env python
"""Batch / HPO entry point: (data-parallel) training of the ATLAS RPV CNN, printing ``FoM: <float>`` for the optimisers.

Flag-compatible with the reference ``train_rpv.py`` (argparse block at :16-32): --input-dir --n-train --n-valid
--n-test --h1..--h4 --dropout --lr --lr-scaling{linear} --optimizer --batch-size --n-epochs --fom{best,last}.
Run one process per GPU, e.g.
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 train_rpv.py --n-epochs 4 --fom best
(a plain ``python train_rpv.py`` is a one-rank world).  Extra flags: --channels (synthetic data), --device, --verbose.
"""
import argparse
import socket
import sys


def parse_args(argv=None):
    p = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    p.add_argument('--input-dir', default='/global/cscratch1/sd/sfarrell/atlas-rpv-images')
    for name, default in (('--n-train', 64000), ('--n-valid', 32000), ('--n-test', 0),
                          ('--h1', 16), ('--h2', 32), ('--h3', 64), ('--h4', 128), ('--batch-size', 64), ('--n-epochs', 4)):
        p.add_argument(name, type=int, default=default)
    p.add_argument('--dropout', type=float, default=0.2)
    p.add_argument('--lr', type=float, default=0.001)
    p.add_argument('--lr-scaling', choices=['linear'])
    p.add_argument('--optimizer', default='Adam')
    p.add_argument('--fom', choices=['best', 'last'])
    p.add_argument('--channels', type=int, default=None, help='channels of the synthetic images when the HDF5 files are absent')
    p.add_argument('--device', default='auto')
    p.add_argument('--verbose', type=int, default=2)
    return p.parse_args(argv)


def main(argv=None):
    import os
    args = parse_args(argv)
    if args.channels:
        os.environ['RPV_CHANNELS'] = str(args.channels)
    from cori_intml_examples_b200.parallel import dist as hvd
    from cori_intml_examples_b200.models.rpv import load_dataset, build_model, train_model

    print('Distributed RPV classifier training')
    hvd.init()
    print('rank %i, local rank %i, size %i, host %s' % (hvd.rank(), hvd.local_rank(), hvd.size(), socket.gethostname()))

    (train_x, train_y, _), (valid_x, valid_y, _), (test_x, test_y, _) = load_dataset(
        args.input_dir, args.n_train, args.n_valid, args.n_test)
    print('train shape:', train_x.shape, 'Mean label:', float(train_y.mean()))
    print('valid shape:', valid_x.shape, 'Mean label:', float(valid_y.mean()))
    if args.n_test > 0:
        print('test shape: ', test_x.shape, 'Mean label:', float(test_y.mean()))

    lr = args.lr * hvd.size() if args.lr_scaling == 'linear' else args.lr
    model = build_model(train_x.shape[1:], conv_sizes=[args.h1, args.h2, args.h3], fc_sizes=[args.h4],
                        dropout=args.dropout, optimizer=args.optimizer, lr=lr, use_horovod=True, device=args.device)
    if hvd.rank() == 0:
        model.summary()

    print('Begin training')
    history = train_model(model, train_input=train_x, train_labels=train_y, valid_input=valid_x, valid_labels=valid_y,
                          batch_size=args.batch_size, n_epochs=args.n_epochs, verbose=args.verbose if hvd.rank() == 0 else 0,
                          use_horovod=True)

    if args.fom == 'best':
        print('FoM:', min(history.history['val_loss']))
    elif args.fom == 'last':
        print('FoM:', history.history['val_loss'][-1])

    if hvd.rank() == 0 and args.n_test > 0:
        score = model.evaluate(test_x, test_y, verbose=0)
        print('Test loss:', score[0])
        print('Test accuracy:', score[1])
    hvd.shutdown()
    return 0


if __name__ == '__main__':
    sys.exit(main())
