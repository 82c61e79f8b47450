"""Flat-name shim: `import ipcluster_magics` registers %ipcluster / %%px when IPython is running."""
from cori_intml_examples_b200.cluster.magics import IPClusterMagics, ipcluster, px, load_ipython_extension, USAGE  # noqa: F401
