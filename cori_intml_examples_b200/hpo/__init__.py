"""Hyper-parameter optimisation: random search over a cluster, grid search + k-fold CV, genetic search over a CLI.

``hpo.Params / hpo.Evaluator / hpo.GeneticOptimizer`` mirror ``crayai.hpo`` as used in CrayHPO_*.ipynb.
"""
from .genetic import Params, Evaluator, GeneticOptimizer
from .random_search import RandomSearch, sample_mnist_space, sample_rpv_space
from .grid import KerasClassifier, GridSearchCV

__all__ = ["Params", "Evaluator", "GeneticOptimizer", "RandomSearch", "sample_mnist_space", "sample_rpv_space",
           "KerasClassifier", "GridSearchCV"]
