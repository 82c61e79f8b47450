"""Layer specifications (Keras 2.2 surface as used by the reference).

Layers here are *descriptions*: shapes, parameter shapes, initialisers and config.  The math lives in the two
backends (``backend_cpu`` = PyTorch-CPU oracle, ``backend_gpu`` = sm_100a engine).  Data format is NHWC
("channels_last", reference mnist.py:30); Conv2D kernels are HWIO and Dense kernels (in, out), i.e. exactly the
Keras checkpoint layouts, so weights move between backends and files without permutation.

Reference call sites: Conv2D/MaxPooling2D/Dropout/Flatten/Dense at mnist.py:48-56 and rpv.py:42-59.
"""
import collections
import math

import numpy as np

_name_counters = collections.defaultdict(int)


def _auto_name(prefix):
    _name_counters[prefix] += 1
    return "%s_%d" % (prefix, _name_counters[prefix])


def reset_name_counters():
    """Restart Keras-style auto naming (conv2d_1, dense_1, ...); handy in tests."""
    _name_counters.clear()


def _pair(v):
    if isinstance(v, (tuple, list)):
        return int(v[0]), int(v[1])
    return int(v), int(v)


class Layer(object):
    """Base class.  ``_inbound`` makes the functional API work: calling a layer on a tensor chains it."""
    _prefix = "layer"
    has_weights = False

    def __init__(self, name=None, input_shape=None):
        self.name = name or _auto_name(self._prefix)
        self.input_shape_arg = tuple(input_shape) if input_shape is not None else None
        self.input_shape = None
        self.output_shape = None
        self._inbound = None
        self.built = False

    # -- functional API: layer(tensor) -> tensor
    def __call__(self, tensor):
        if not isinstance(tensor, SymbolicTensor):
            raise TypeError("layers are called on tensors produced by Input(...) or by other layers")
        out = SymbolicTensor(self, tensor)
        return out

    def build(self, input_shape):
        self.input_shape = tuple(input_shape)
        self.output_shape = self.compute_output_shape(self.input_shape)
        self.built = True

    def compute_output_shape(self, input_shape):
        return tuple(input_shape)

    def weight_shapes(self):
        return []

    def weight_names(self):
        return []

    def init_weights(self, rng):
        return []

    def count_params(self):
        return int(sum(int(np.prod(s)) for s in self.weight_shapes()))

    def get_config(self):
        return {"name": self.name, "trainable": True}

    @property
    def class_name(self):
        return type(self).__name__


class SymbolicTensor(object):
    """Node of the (linear) functional graph: remembers the producing layer and its input tensor."""

    def __init__(self, layer, parent):
        self.layer = layer
        self.parent = parent

    def chain(self):
        layers, t = [], self
        while t is not None:
            layers.append(t.layer)
            t = t.parent
        return layers[::-1]


class InputLayer(Layer):
    _prefix = "input"

    def __init__(self, shape, name=None):
        super(InputLayer, self).__init__(name=name)
        self.shape = tuple(int(s) for s in shape)

    def compute_output_shape(self, input_shape):
        return self.shape

    def get_config(self):
        return {"batch_input_shape": [None] + list(self.shape), "dtype": "float32", "sparse": False, "name": self.name}


def Input(shape, name=None):
    """Functional-API entry point (reference rpv.py:42)."""
    layer = InputLayer(shape, name=name)
    layer.build(shape)
    return SymbolicTensor(layer, None)


def glorot_uniform(shape, fan_in, fan_out, rng):
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-limit, limit, size=shape).astype(np.float32)


class Conv2D(Layer):
    _prefix = "conv2d"
    has_weights = True

    def __init__(self, filters, kernel_size=(3, 3), strides=(1, 1), padding="valid", activation=None,
                 name=None, input_shape=None, **kwargs):
        super(Conv2D, self).__init__(name=name, input_shape=input_shape)
        self.filters = int(filters)
        self.kernel_size = _pair(kernel_size)
        self.strides = _pair(strides)
        if padding not in ("valid", "same"):
            raise ValueError("padding must be 'valid' or 'same'")
        self.padding = padding
        self.activation = activation or "linear"

    def pad_amounts(self, size, k, s):
        """(before, after) padding along one axis, TensorFlow 'SAME' rule."""
        if self.padding == "valid":
            return 0, 0
        out = -(-size // s)
        total = max((out - 1) * s + k - size, 0)
        return total // 2, total - total // 2

    def compute_output_shape(self, input_shape):
        h, w, _ = input_shape
        kh, kw = self.kernel_size
        sh, sw = self.strides
        if self.padding == "same":
            return (-(-h // sh), -(-w // sw), self.filters)
        return ((h - kh) // sh + 1, (w - kw) // sw + 1, self.filters)

    def weight_shapes(self):
        cin = self.input_shape[-1]
        return [self.kernel_size + (cin, self.filters), (self.filters,)]

    def weight_names(self):
        return [self.name + "/kernel:0", self.name + "/bias:0"]

    def init_weights(self, rng):
        kh, kw = self.kernel_size
        cin = self.input_shape[-1]
        k = glorot_uniform(self.weight_shapes()[0], kh * kw * cin, kh * kw * self.filters, rng)
        return [k, np.zeros((self.filters,), np.float32)]

    def get_config(self):
        cfg = {"name": self.name, "trainable": True, "filters": self.filters, "kernel_size": list(self.kernel_size),
               "strides": list(self.strides), "padding": self.padding, "data_format": "channels_last",
               "dilation_rate": [1, 1], "activation": self.activation, "use_bias": True,
               "kernel_initializer": {"class_name": "VarianceScaling",
                                      "config": {"scale": 1.0, "mode": "fan_avg", "distribution": "uniform", "seed": None}},
               "bias_initializer": {"class_name": "Zeros", "config": {}},
               "kernel_regularizer": None, "bias_regularizer": None, "activity_regularizer": None,
               "kernel_constraint": None, "bias_constraint": None}
        if self.input_shape_arg is not None:
            cfg["batch_input_shape"] = [None] + list(self.input_shape_arg)
            cfg["dtype"] = "float32"
        return cfg


class MaxPooling2D(Layer):
    _prefix = "max_pooling2d"

    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", name=None, **kwargs):
        super(MaxPooling2D, self).__init__(name=name)
        self.pool_size = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool_size
        if padding != "valid":
            raise NotImplementedError("only 'valid' pooling is used by the reference")
        self.padding = padding

    def compute_output_shape(self, input_shape):
        h, w, c = input_shape
        return ((h - self.pool_size[0]) // self.strides[0] + 1, (w - self.pool_size[1]) // self.strides[1] + 1, c)

    def get_config(self):
        return {"name": self.name, "trainable": True, "pool_size": list(self.pool_size), "padding": self.padding,
                "strides": list(self.strides), "data_format": "channels_last"}


class Dropout(Layer):
    _prefix = "dropout"

    def __init__(self, rate, name=None, **kwargs):
        super(Dropout, self).__init__(name=name)
        self.rate = float(rate)
        if not 0.0 <= self.rate < 1.0:
            raise ValueError("dropout rate must be in [0, 1)")

    def get_config(self):
        return {"name": self.name, "trainable": True, "rate": self.rate, "noise_shape": None, "seed": None}


class Flatten(Layer):
    _prefix = "flatten"

    def compute_output_shape(self, input_shape):
        return (int(np.prod(input_shape)),)

    def get_config(self):
        return {"name": self.name, "trainable": True, "data_format": "channels_last"}


class Dense(Layer):
    _prefix = "dense"
    has_weights = True

    def __init__(self, units, activation=None, name=None, input_shape=None, **kwargs):
        super(Dense, self).__init__(name=name, input_shape=input_shape)
        self.units = int(units)
        self.activation = activation or "linear"

    def compute_output_shape(self, input_shape):
        return tuple(input_shape[:-1]) + (self.units,)

    def weight_shapes(self):
        return [(self.input_shape[-1], self.units), (self.units,)]

    def weight_names(self):
        return [self.name + "/kernel:0", self.name + "/bias:0"]

    def init_weights(self, rng):
        fin = self.input_shape[-1]
        return [glorot_uniform((fin, self.units), fin, self.units, rng), np.zeros((self.units,), np.float32)]

    def get_config(self):
        cfg = {"name": self.name, "trainable": True, "units": self.units, "activation": self.activation, "use_bias": True,
               "kernel_initializer": {"class_name": "VarianceScaling",
                                      "config": {"scale": 1.0, "mode": "fan_avg", "distribution": "uniform", "seed": None}},
               "bias_initializer": {"class_name": "Zeros", "config": {}},
               "kernel_regularizer": None, "bias_regularizer": None, "activity_regularizer": None,
               "kernel_constraint": None, "bias_constraint": None}
        if self.input_shape_arg is not None:
            cfg["batch_input_shape"] = [None] + list(self.input_shape_arg)
            cfg["dtype"] = "float32"
        return cfg


LAYER_CLASSES = {c.__name__: c for c in (InputLayer, Conv2D, MaxPooling2D, Dropout, Flatten, Dense)}


def layer_from_config(class_name, config):
    """Rebuild a layer from a Keras ``model_config`` entry (used by load_model)."""
    cfg = dict(config)
    name = cfg.get("name")
    bis = cfg.get("batch_input_shape")
    ishape = tuple(bis[1:]) if bis else None
    if class_name == "InputLayer":
        return InputLayer(ishape, name=name)
    if class_name == "Conv2D":
        return Conv2D(cfg["filters"], tuple(cfg["kernel_size"]), strides=tuple(cfg.get("strides", (1, 1))),
                      padding=cfg.get("padding", "valid"), activation=cfg.get("activation"), name=name, input_shape=ishape)
    if class_name == "MaxPooling2D":
        return MaxPooling2D(tuple(cfg.get("pool_size", (2, 2))), strides=tuple(cfg["strides"]) if cfg.get("strides") else None, name=name)
    if class_name == "Dropout":
        return Dropout(cfg["rate"], name=name)
    if class_name == "Flatten":
        return Flatten(name=name)
    if class_name == "Dense":
        return Dense(cfg["units"], activation=cfg.get("activation"), name=name, input_shape=ishape)
    raise ValueError("unsupported layer class in checkpoint: %s" % class_name)
