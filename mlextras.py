"""Flat-name shim: `from mlextras import IPyParallelLogger, configure_session`."""
from cori_intml_examples_b200.mlextras import IPyParallelLogger, configure_session, SessionConfig  # noqa: F401
