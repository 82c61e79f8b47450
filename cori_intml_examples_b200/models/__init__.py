"""Model builders and data loaders with the reference's public surface (mnist.py, rpv.py)."""
from . import mnist, rpv
