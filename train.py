#!/usr/bin/env python
"""Alias of train_mnist.py under the file name CrayHPO_mnist.ipynb:75 evaluates (`python train.py --epochs N --h1 .. --dropout ..`);
the reference never shipped that file.  Prints ``FoM: <val_loss>``."""
import sys

from train_mnist import main

if __name__ == '__main__':
    sys.exit(main())
