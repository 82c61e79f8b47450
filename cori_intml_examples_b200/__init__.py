"""cori_intml_examples_b200 -- interactive distributed deep learning on 8xB200, Blackwell-native.

Same workflows as mlhenderson/cori-intml-examples (Keras-style CNN builders, Horovod-style data
parallelism, IPyParallel-style cluster + load-balanced HPO, live HPO monitoring), rebuilt on
PyTorch tensors + hand-written sm_100a CUDA kernels (tcgen05/TMEM/TMA) + peer-to-peer NVLink
collectives.  See DESIGN.md for the layer map.
"""
__version__ = "0.1.0"
