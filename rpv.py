"""Flat-name shim so reference notebooks keep working: `import rpv` == cori_intml_examples_b200.models.rpv."""
from cori_intml_examples_b200.models.rpv import *  # noqa: F401,F403
from cori_intml_examples_b200.models.rpv import build_model  # noqa: F401
from cori_intml_examples_b200.models.rpv import load_file, load_dataset, train_model, build_legacy_model  # noqa: F401,E402
