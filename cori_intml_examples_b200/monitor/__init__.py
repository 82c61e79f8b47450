"""Live HPO monitoring (headless state + text/HTML/SVG renderers) and post-hoc analysis metrics."""
from .hpo_widgets import ModelPlot, ParamSpanWidget, ModelController, ModelTaskData, ModelPlotTable
from . import analysis

__all__ = ["ModelPlot", "ParamSpanWidget", "ModelController", "ModelTaskData", "ModelPlotTable", "analysis"]
