"""Keras-style numerical engine (layers, models, optimizers, callbacks) with a CPU oracle and a B200 backend."""
from . import layers, optimizers, callbacks
from .layers import Input, Conv2D, MaxPooling2D, Dropout, Flatten, Dense
from .model import Model, Sequential
from .saving import load_model, save_model

__all__ = ["layers", "optimizers", "callbacks", "Input", "Conv2D", "MaxPooling2D", "Dropout", "Flatten", "Dense",
           "Model", "Sequential", "load_model", "save_model"]
