"""Flat-name shim: `from hpo_widgets import ParamSpanWidget, ModelPlot, ...`."""
from cori_intml_examples_b200.monitor.hpo_widgets import (ModelPlot, ParamSpanWidget, ModelController,  # noqa: F401
                                                          ModelTaskData, ModelPlotTable)
