"""Flat-name shim so reference notebooks keep working: `import mnist` == cori_intml_examples_b200.models.mnist."""
from cori_intml_examples_b200.models.mnist import *  # noqa: F401,F403
from cori_intml_examples_b200.models.mnist import build_model  # noqa: F401
from cori_intml_examples_b200.models.mnist import load_data, n_classes, img_rows, img_cols  # noqa: F401,E402
