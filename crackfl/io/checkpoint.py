"""Native checkpoint: flat fp32 arenas + JSON manifest (SURVEY §5.4 "New framework (c)").

The reference persists only ``./server_weights/weights.pickle`` (``fl_server.py:104-105``,
never read back) and no round counter, so a restarted server starts again from round 0
with a fresh model.  This format lets the server resume: the manifest carries round,
model_version, the client registry and the server-optimiser (FedAdam m/v) state; each
arena is one little-endian float32 file that can be mapped straight into the device
arena.  Every file is written tmp + rename, and the checkpoint AS A WHOLE commits at one
point: arenas are written under names that carry a save id (``global.<id>.f32``), the
manifest that names them is renamed into place last, and only then are the previous
save's arena files deleted.  A crash at any moment leaves either the old manifest with
its old arenas or the new manifest with its new arenas, never new weights under the old
round number.
"""
from __future__ import annotations

import json
import os
import time
from typing import Dict, List, Optional, Sequence

import numpy as np

from .weights import flatten, unflatten

FORMAT = "crackfl-ckpt-1"


def _atomic_write(path: str, data: bytes):
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        f.write(data)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)


def save_checkpoint(dirpath: str, spec, weights: Sequence[np.ndarray], meta: Optional[dict] = None,
                    extra_arenas: Optional[Dict[str, Sequence[np.ndarray]]] = None) -> str:
    os.makedirs(dirpath, exist_ok=True)
    arenas = {"global": weights}
    arenas.update(extra_arenas or {})
    files = {}
    old_files = []
    try:
        with open(os.path.join(dirpath, "manifest.json")) as f:
            old_files = list(json.load(f).get("arenas", {}).values())
    except (OSError, ValueError):
        pass
    mv = (meta or {}).get("model_version")
    save_id = f"{mv if mv is not None else 'x'}-{time.time_ns():x}"
    for name, wl in arenas.items():
        if wl is None:
            continue
        fn = f"{name}.{save_id}.f32"
        _atomic_write(os.path.join(dirpath, fn), flatten(wl, spec).astype("<f4").tobytes())
        files[name] = fn
    manifest = {
        "format": FORMAT,
        "saved_at": time.time(),
        "arena_numel": spec.arena_numel,
        "num_params": spec.num_params,
        "tensors": [{"name": t.name, "shape": list(t.shape), "offset": t.offset, "trainable": t.trainable}
                    for t in spec.tensors],
        "arenas": files,
        "meta": meta or {},
    }
    _atomic_write(os.path.join(dirpath, "manifest.json"), json.dumps(manifest, indent=1).encode())  # the commit point
    for fn in old_files:
        if fn not in files.values():
            try:
                os.remove(os.path.join(dirpath, fn))
            except OSError:
                pass
    return os.path.join(dirpath, "manifest.json")


def load_checkpoint(dirpath: str, spec) -> dict:
    with open(os.path.join(dirpath, "manifest.json")) as f:
        man = json.load(f)
    if man.get("format") != FORMAT:
        raise ValueError(f"unknown checkpoint format {man.get('format')!r}")
    names = [t["name"] for t in man["tensors"]]
    if names != spec.names():
        raise ValueError("checkpoint tensor table does not match the model spec")
    out = {"meta": man.get("meta", {}), "arenas": {}}
    for name, fn in man["arenas"].items():
        flat = np.fromfile(os.path.join(dirpath, fn), dtype="<f4")
        if flat.size != man["arena_numel"]:
            raise ValueError(f"{fn}: expected {man['arena_numel']} floats, found {flat.size}")
        out["arenas"][name] = unflatten(flat, spec)
    out["weights"] = out["arenas"]["global"]
    return out


def has_checkpoint(dirpath: str) -> bool:
    return os.path.exists(os.path.join(dirpath, "manifest.json"))
