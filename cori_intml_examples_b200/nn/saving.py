"""Keras 2.2 full-model HDF5 checkpoints: ``model.save`` / ``load_model`` / ``save_weights`` / ``load_weights``.

File layout: root attrs ``keras_version, backend, model_config, training_config``;
``/model_weights`` with attr ``layer_names`` and one group per layer carrying attr ``weight_names`` and datasets
named by those (Conv2D kernel HWIO, Dense kernel (in,out) with rows in (H,W,C) flatten order);
``/optimizer_weights`` with attr ``weight_names`` and ``iterations`` + slot tensors in Keras ``optimizer.weights`` order.
Reference call sites: ModelCheckpoint rpv.py:100-101, DistHPO_mnist.ipynb:184-185; load_model DistHPO_mnist.ipynb:542.
"""
import json

import numpy as np

from ..utils import hdf5
from . import layers as L
from . import optimizers as optimizers_mod

KERAS_VERSION = b"2.2.4"
BACKEND = b"tensorflow"

_SLOT_NAMES = {"SGD": ["moments"], "Adam": ["m", "v"], "Nadam": ["m", "v"], "Adadelta": ["accumulator", "delta_accumulator"]}


def _bytes_list(names):
    if not names:
        return np.zeros((0,), dtype="S1")
    return np.array([n.encode("utf8") for n in names])


def _write_model_weights(group, model, weights):
    group.attrs["layer_names"] = _bytes_list([l.name for l in model.layers])
    group.attrs["backend"] = BACKEND
    group.attrs["keras_version"] = KERAS_VERSION
    it = iter(weights)
    for layer in model.layers:
        g = group.create_group(layer.name)
        names = layer.weight_names()
        g.attrs["weight_names"] = _bytes_list(names)
        for nm in names:
            g.create_dataset(nm, data=np.asarray(next(it), np.float32))


def save_weights(model, filepath):
    with hdf5.File(filepath, "w") as f:
        _write_model_weights(f, model, model.get_weights())


def save_model(model, filepath, include_optimizer=True):
    weights = model.get_weights()
    with hdf5.File(filepath, "w") as f:
        f.attrs["keras_version"] = KERAS_VERSION
        f.attrs["backend"] = BACKEND
        cls = "Sequential" if type(model).__name__ == "Sequential" else "Model"
        f.attrs["model_config"] = json.dumps({"class_name": cls, "config": model.get_config()}).encode("utf8")
        _write_model_weights(f.create_group("model_weights"), model, weights)
        if include_optimizer and model.optimizer is not None:
            opt = model.optimizer
            model.backend.sync_optimizer_to_host()
            f.attrs["training_config"] = json.dumps({
                "optimizer_config": {"class_name": opt.class_name, "config": opt.get_config()},
                "loss": model.loss, "metrics": ["accuracy"] if len(model.metrics_names) > 1 else [],
                "sample_weight_mode": None, "loss_weights": None}).encode("utf8")
            slots = model.backend.get_slots()
            arrays = opt.state_arrays(slots)
            prefix = "training/" + opt.class_name + "/"
            names = ["iterations:0"]
            if opt.class_name == "Nadam":
                names.append("m_schedule:0")
            for sname, s in zip(_SLOT_NAMES[opt.class_name], slots):
                names += ["%s_%d:0" % (sname, i) for i in range(len(s))]
            names = [prefix + n for n in names]
            g = f.create_group("optimizer_weights")
            g.attrs["weight_names"] = _bytes_list(names)
            for nm, arr in zip(names, arrays):
                g.create_dataset(nm, data=arr)


def _decode(v):
    if isinstance(v, np.ndarray) and v.shape == ():
        v = v[()]
    if isinstance(v, (bytes, np.bytes_)):
        return v.decode("utf8")
    return v


def _read_layer_weights(group):
    names = [_decode(n) for n in np.atleast_1d(group.attrs["layer_names"])] if len(np.atleast_1d(group.attrs["layer_names"])) else []
    out = []
    for lname in names:
        g = group[lname]
        wn = np.atleast_1d(g.attrs["weight_names"])
        out.append((lname, [np.asarray(g[_decode(n)][...], np.float32) for n in wn if len(n)]))
    return out


def load_weights(model, filepath):
    with hdf5.File(filepath, "r") as f:
        g = f["model_weights"] if "model_weights" in f else f
        per_layer = _read_layer_weights(g)
    # key on order, not on literal names (auto-name counters differ between processes)
    flat = [w for _, ws in per_layer for w in ws]
    model.set_weights(flat)


def model_from_config(cfg):
    from .model import Model, Sequential
    cls, conf = cfg["class_name"], cfg["config"]
    layer_cfgs = conf["layers"] if isinstance(conf, dict) else conf
    layers = [L.layer_from_config(lc["class_name"], lc["config"]) for lc in layer_cfgs]
    name = conf.get("name") if isinstance(conf, dict) else None
    if cls == "Sequential":
        return Sequential(layers, name=name)
    if not isinstance(layers[0], L.InputLayer):
        raise ValueError("functional model config must start with an InputLayer")
    layers[0].build(layers[0].shape)
    return Model(name=name, _layers=layers)


def load_model(filepath, compile=True, device="auto"):
    """Rebuild a model from the HDF5 file alone (architecture + weights [+ optimizer state])."""
    with hdf5.File(filepath, "r") as f:
        cfg = f.attrs.get("model_config")
        if cfg is None:
            raise ValueError("no model_config in %s (weights-only file?)" % filepath)
        model = model_from_config(json.loads(_decode(cfg)))
        per_layer = _read_layer_weights(f["model_weights"])
        tc = f.attrs.get("training_config")
        opt_arrays = None
        if compile and tc is not None and "optimizer_weights" in f:
            g = f["optimizer_weights"]
            opt_arrays = [np.asarray(g[_decode(n)][...]) for n in np.atleast_1d(g.attrs["weight_names"])]
    flat = [w for _, ws in per_layer for w in ws]
    if compile and tc is not None:
        tc = json.loads(_decode(tc))
        oc = tc["optimizer_config"]
        opt = optimizers_mod.get(oc["class_name"], **{k: v for k, v in oc["config"].items()
                                                      if k in ("lr", "beta_1", "beta_2", "epsilon", "decay", "rho", "momentum", "schedule_decay")
                                                      and v is not None})
        model.compile(optimizer=opt, loss=tc["loss"], metrics=tc.get("metrics") or None, device=device)
        model.set_weights(flat)
        if opt_arrays is not None:
            slots = [[np.zeros_like(w) for w in flat] for _ in range(opt.n_slots)]
            try:
                opt.load_state_arrays(opt_arrays, slots)
                model.backend.set_slots(slots)
                if hasattr(model.backend, "push_hyper"):
                    model.backend.push_hyper()
            except (IndexError, ValueError):
                pass        # optimizer state of a different shape: keep fresh slots
    else:
        model._device_pref = device
        model.set_weights(flat)
    return model
