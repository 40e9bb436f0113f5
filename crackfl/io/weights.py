"""Weight-list serialisation: reference pickle format, safe raw framing, flat arena views.

The reference ships model weights as ``pickle.dumps(model.get_weights())`` — a Python
``list`` of 112 float32 ``numpy.ndarray`` — both on the wire (``fl_client.py:63``,
``fl_server.py:24,179``) and on disk (``client_fit_model.py:239-240``,
``fl_server.py:104-105``).  ``pickle.loads`` on network bytes is arbitrary code
execution, so:

* ``loads_pickle`` uses a restricted unpickler that only admits the globals a list of
  ndarrays needs; anything else raises ``UnsafePickleError``;
* ``dumps_raw`` / ``loads_raw`` is the default framing of this framework: a small header
  plus the raw little-endian tensor bytes.  ``loads_auto`` sniffs which one it got, and
  only takes the pickle branch when ``allow_pickle`` is set.
"""
from __future__ import annotations

import io
import pickle
import struct
from typing import List, Sequence

import numpy as np

RAW_MAGIC = b"CKFLRAW1"
_DTYPES = {0: np.float32, 1: np.float16, 2: np.uint16, 3: np.float64, 4: np.int32, 5: np.uint8, 6: np.int64}
_DTYPE_CODES = {np.dtype(v).str: k for k, v in _DTYPES.items()}


class UnsafePickleError(pickle.UnpicklingError):
    pass


_ALLOWED = {
    ("numpy.core.multiarray", "_reconstruct"), ("numpy._core.multiarray", "_reconstruct"),
    ("numpy.core.multiarray", "scalar"), ("numpy._core.multiarray", "scalar"),
    ("numpy", "ndarray"), ("numpy", "dtype"),
    ("numpy.core.numeric", "_frombuffer"), ("numpy._core.numeric", "_frombuffer"),
    ("builtins", "list"), ("builtins", "tuple"), ("builtins", "dict"),
    ("collections", "OrderedDict"),
}


class _RestrictedUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if (module, name) in _ALLOWED:
            return super().find_class(module, name)
        raise UnsafePickleError(f"refusing to unpickle global {module}.{name}")


def dumps_pickle(weights: Sequence[np.ndarray], protocol: int = 4) -> bytes:
    """Exactly what the reference sends: ``pickle.dumps(list_of_ndarrays)``."""
    return pickle.dumps([np.asarray(w) for w in weights], protocol=protocol)


def loads_pickle(buf: bytes) -> List[np.ndarray]:
    obj = _RestrictedUnpickler(io.BytesIO(buf)).load()
    if not isinstance(obj, (list, tuple)) or not all(isinstance(a, np.ndarray) for a in obj):
        raise UnsafePickleError("payload is not a list of ndarrays")
    if any(a.dtype.hasobject for a in obj):
        raise UnsafePickleError("object arrays are not accepted")
    return list(obj)


def dumps_raw(weights: Sequence[np.ndarray]) -> bytes:
    """``CKFLRAW1 | u32 count | per tensor: u8 dtype, u8 ndim, u32 dims[ndim] | payload...``"""
    hdr = [RAW_MAGIC, struct.pack("<I", len(weights))]
    body = []
    for w in weights:
        a = np.asarray(w)
        if not a.flags.c_contiguous:  # (np.ascontiguousarray would also turn a 0-d array into shape (1,))
            a = np.ascontiguousarray(a)
        if a.dtype.byteorder == ">":
            a = a.astype(a.dtype.newbyteorder("<"))
        code = _DTYPE_CODES.get(a.dtype.str)
        if code is None:
            raise TypeError(f"unsupported dtype {a.dtype}")
        hdr.append(struct.pack("<BB", code, a.ndim) + struct.pack(f"<{a.ndim}I", *a.shape))
        body.append(a.tobytes())
    return b"".join(hdr + body)


def loads_raw(buf: bytes) -> List[np.ndarray]:
    mv = memoryview(buf)
    if bytes(mv[:8]) != RAW_MAGIC:
        raise ValueError("not a CKFLRAW1 payload")
    (count,) = struct.unpack_from("<I", mv, 8)
    pos = 12
    metas = []
    for _ in range(count):
        code, ndim = struct.unpack_from("<BB", mv, pos)
        pos += 2
        shape = struct.unpack_from(f"<{ndim}I", mv, pos)
        pos += 4 * ndim
        if code not in _DTYPES:
            raise ValueError("bad dtype code")
        metas.append((np.dtype(_DTYPES[code]), shape))
    out = []
    for dt, shape in metas:
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        nb = n * dt.itemsize
        if pos + nb > len(mv):
            raise ValueError("truncated payload")
        out.append(np.frombuffer(mv[pos:pos + nb], dtype=dt).reshape(shape).copy())
        pos += nb
    return out


def loads_auto(buf: bytes, allow_pickle: bool = False) -> List[np.ndarray]:
    if bytes(buf[:8]) == RAW_MAGIC:
        return loads_raw(buf)
    if not allow_pickle:
        raise UnsafePickleError("pickle payloads are disabled (pass allow_pickle=True / --compat-pickle)")
    return loads_pickle(buf)


def dumps(weights: Sequence[np.ndarray], fmt: str = "raw") -> bytes:
    if fmt == "raw":
        return dumps_raw(weights)
    if fmt == "pickle":
        return dumps_pickle(weights)
    raise ValueError(fmt)


def save_pickle_file(path: str, weights: Sequence[np.ndarray]):
    """``./saved_weight/weights.pickle`` / ``./server_weights/weights.pickle`` equivalent;
    creates the directory (the reference does not: quirk Q13)."""
    import os
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(dumps_pickle(weights))


def load_pickle_file(path: str) -> List[np.ndarray]:
    with open(path, "rb") as f:
        return loads_pickle(f.read())


# -- flat arena <-> list ---------------------------------------------------------------
def flatten(weights: Sequence[np.ndarray], spec) -> np.ndarray:
    """Concatenate into the flat fp32 arena layout (zero padded to ``spec.arena_numel``)."""
    if len(weights) != spec.num_tensors:
        raise ValueError(f"expected {spec.num_tensors} tensors, got {len(weights)}")
    flat = np.zeros((spec.arena_numel,), np.float32)
    for t, w in zip(spec.tensors, weights):
        a = np.asarray(w, np.float32)
        if tuple(a.shape) != t.shape:
            raise ValueError(f"{t.name}: expected shape {t.shape}, got {tuple(a.shape)}")
        flat[t.offset:t.offset + t.numel] = a.reshape(-1)
    return flat


def unflatten(flat: np.ndarray, spec) -> List[np.ndarray]:
    flat = np.asarray(flat)
    return [flat[t.offset:t.offset + t.numel].reshape(t.shape).astype(np.float32, copy=True) for t in spec.tensors]
