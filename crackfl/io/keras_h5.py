"""Keras-layout HDF5 weight files: ``crack_segmentation.h5`` compatibility (reference C21/C22).

``test/Segmentation.py:177-179`` writes ``crack_segmentation.h5`` with
``ModelCheckpoint(save_best_only=True)``: a Keras 2.x *full-model* file:

    /                      attrs: keras_version, backend, model_config (JSON), training_config (JSON)
    /model_weights         attrs: layer_names [S], backend, keras_version
    /model_weights/<layer> attrs: weight_names [S]   e.g. b"conv2d/kernel:0"
    /model_weights/<layer>/<layer>/<weight>:0        float32 datasets
    /optimizer_weights     (Adam slots; optional)

``model.save_weights("x.h5")`` writes the content of ``/model_weights`` at the root instead.  Both are
accepted on load.  Loading is NAME-KEYED, never positional: Keras orders ``get_weights()`` by graph depth
with ties the reference code cannot pin down, and layer suffixes shift when several models
were built in one session (``conv2d_9`` instead of ``conv2d``), so layers are matched per class in order of
their numeric suffix and tensors by weight name, with shape checks.

STATUS: self-consistency only (no file written by real Keras is available; h5py/TensorFlow cannot be
installed here).  SavedModel directories (``model.save("my_model")``, ``test/Segmentation.py:187``) need
TensorFlow and are out of scope; the native checkpoint (io/checkpoint.py) covers that role.
"""
from __future__ import annotations

import json
import re
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .hdf5 import H5Dataset, H5Group, read_h5, write_h5

KERAS_VERSION = "2.11.0"
BACKEND = "tensorflow"


def keras_layer_sequence(spec) -> List[Tuple[str, str, dict, List[str]]]:
    """All layers of the functional model in construction order: (name, class, config, inbound layer names)."""
    out: List[Tuple[str, str, dict, List[str]]] = []
    counters: Dict[str, int] = {}

    def nm(base):
        k = counters.get(base, 0)
        counters[base] = k + 1
        return base if k == 0 else f"{base}_{k}"

    def add(base, cls, cfg, inbound):
        n = nm(base)
        out.append((n, cls, dict(cfg, name=n), inbound))
        return n

    x = add("input_1", "InputLayer", {"batch_input_shape": [None, None, None, spec.in_channels], "dtype": "float32"}, [])
    conv = lambda f, k, s, i: add("conv2d", "Conv2D", {"filters": f, "kernel_size": [k, k], "strides": [s, s],
                                                      "padding": "same", "use_bias": True}, [i])
    bn = lambda i: add("batch_normalization", "BatchNormalization", {"momentum": spec.bn_momentum, "epsilon": spec.bn_eps,
                                                                    "axis": [3]}, [i])
    act = lambda i, a="relu": add("activation", "Activation", {"activation": a}, [i])
    x = act(bn(conv(spec.entry, 3, 2, x)))
    prev = x
    for f in spec.down:
        for _ in range(2):
            x = act(x)
            x = add("separable_conv2d", "SeparableConv2D", {"filters": f, "kernel_size": [3, 3], "padding": "same"}, [x])
            x = bn(x)
        x = add("max_pooling2d", "MaxPooling2D", {"pool_size": [3, 3], "strides": [2, 2], "padding": "same"}, [x])
        r = conv(f, 1, 2, prev)
        x = add("add", "Add", {}, [x, r])
        prev = x
    for f in spec.up:
        for _ in range(2):
            x = act(x)
            x = add("conv2d_transpose", "Conv2DTranspose", {"filters": f, "kernel_size": [3, 3], "padding": "same"}, [x])
            x = bn(x)
        x = add("up_sampling2d", "UpSampling2D", {"size": [2, 2]}, [x])
        r = add("up_sampling2d", "UpSampling2D", {"size": [2, 2]}, [prev])
        r = conv(f, 1, 1, r)
        x = add("add", "Add", {}, [x, r])
        prev = x
    add("conv2d", "Conv2D", {"filters": 1, "kernel_size": [1, 1], "padding": "same", "activation": "sigmoid"}, [x])
    return out


def model_config_json(spec, input_hw=(128, 128)) -> str:
    layers = []
    seq = keras_layer_sequence(spec)
    for name, cls, cfg, inbound in seq:
        cfg = dict(cfg)
        if cls == "InputLayer":
            cfg["batch_input_shape"] = [None, input_hw[0], input_hw[1], spec.in_channels]
        layers.append({"class_name": cls, "config": cfg, "name": name,
                       "inbound_nodes": [[[i, 0, 0, {}] for i in inbound]] if inbound else []})
    return json.dumps({"class_name": "Functional", "config": {
        "name": "model", "layers": layers, "input_layers": [[seq[0][0], 0, 0]], "output_layers": [[seq[-1][0], 0, 0]]}})


TRAINING_CONFIG = json.dumps({"loss": "binary_crossentropy", "metrics": [[{"class_name": "MeanMetricWrapper",
                              "config": {"name": "accuracy", "dtype": "float32", "fn": "binary_accuracy"}}]],
                              "optimizer_config": {"class_name": "Adam", "config": {
                                  "name": "Adam", "learning_rate": 0.001, "beta_1": 0.9, "beta_2": 0.999,
                                  "epsilon": 1e-07, "amsgrad": False}}})


def save_keras_h5(path: str, spec, weights: Sequence[np.ndarray], full_model: bool = True, input_hw=(128, 128)):
    """Write ``weights`` (UNetSpec order) as a Keras full-model (.h5 of ModelCheckpoint) or weights-only file."""
    if len(weights) != spec.num_tensors:
        raise ValueError(f"expected {spec.num_tensors} tensors")
    root = H5Group()
    wroot = root.require_group("model_weights") if full_model else root
    by_layer: Dict[str, List[Tuple[str, np.ndarray]]] = {}
    for t, w in zip(spec.tensors, weights):
        by_layer.setdefault(t.layer, []).append((t.name, np.asarray(w, np.float32).reshape(t.shape)))
    names = [n for n, *_ in keras_layer_sequence(spec)]
    for lname in names:
        g = wroot.require_group(lname)
        ws = by_layer.get(lname, [])
        g.attrs["weight_names"] = [n.encode() for n, _ in ws]
        for n, arr in ws:
            parts = n.split("/")
            sub = g.require_group("/".join(parts[:-1]))
            sub.children[parts[-1]] = H5Dataset(arr)
    wroot.attrs["layer_names"] = [n.encode() for n in names]
    wroot.attrs["backend"] = BACKEND
    wroot.attrs["keras_version"] = KERAS_VERSION
    if full_model:
        root.attrs["keras_version"] = KERAS_VERSION
        root.attrs["backend"] = BACKEND
        root.attrs["model_config"] = model_config_json(spec, input_hw)
        root.attrs["training_config"] = TRAINING_CONFIG
    write_h5(path, root)


def _as_str_list(v) -> List[str]:
    if v is None:
        return []
    arr = np.atleast_1d(np.asarray(v, dtype=object) if not isinstance(v, np.ndarray) else v)
    return [x.decode("utf8") if isinstance(x, (bytes, np.bytes_)) else str(x) for x in arr.tolist()]


def _layer_class(name: str) -> Tuple[str, int]:
    m = re.match(r"^(.*?)(?:_(\d+))?$", name)
    return m.group(1), int(m.group(2) or 0)


def load_keras_h5(path: str, spec, strict: bool = True) -> List[np.ndarray]:
    """Read a Keras .h5 (full-model or weights-only) into UNetSpec order, matching by layer class + weight name."""
    root = read_h5(path)
    wroot = root["model_weights"] if "model_weights" in root else root
    layer_names = _as_str_list(wroot.attrs.get("layer_names"))
    if not layer_names:
        layer_names = [k for k, v in wroot.children.items() if isinstance(v, H5Group)]
    file_layers: Dict[str, Dict[str, np.ndarray]] = {}
    for ln in layer_names:
        if ln not in wroot.children:
            continue
        g = wroot.children[ln]
        tensors = {p.split("/")[-1].split(":")[0]: d.data for p, d in g.visit_datasets()}
        if tensors:
            file_layers[ln] = tensors
    # group by class, ordered by numeric suffix (construction order inside a Keras session)
    by_class: Dict[str, List[str]] = {}
    for ln in file_layers:
        by_class.setdefault(_layer_class(ln)[0], []).append(ln)
    for c in by_class:
        by_class[c].sort(key=lambda n: _layer_class(n)[1])
    ours: Dict[str, List[str]] = {}
    for l in spec.layers:
        ours.setdefault(_layer_class(l.name)[0], []).append(l.name)
    mapping: Dict[str, str] = {}
    for c, lst in ours.items():
        theirs = by_class.get(c, [])
        if len(theirs) < len(lst):
            if strict:
                raise KeyError(f"file has {len(theirs)} '{c}' layers with weights, model needs {len(lst)}")
        for a, b in zip(lst, theirs):
            mapping[a] = b
    out = []
    for t in spec.tensors:
        src = mapping.get(t.layer)
        arr = file_layers.get(src, {}).get(t.weight) if src else None
        if arr is None:
            raise KeyError(f"weight {t.name} not found in {path} (mapped layer: {src})")
        if tuple(arr.shape) != t.shape:
            raise ValueError(f"{t.name}: file has shape {tuple(arr.shape)}, model expects {t.shape}")
        out.append(np.ascontiguousarray(arr, dtype=np.float32))
    return out
