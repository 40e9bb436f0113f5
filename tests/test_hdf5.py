"""Pure-Python HDF5 subset + Keras-layout weight files (SURVEY §4 'Round-trip', hard part H2)."""
import struct

import numpy as np
import pytest

from crackfl.io import hdf5 as h5
from crackfl.io.keras_h5 import keras_layer_sequence, load_keras_h5, model_config_json, save_keras_h5
from crackfl.models.spec import REFERENCE_SPEC as S


def test_generic_roundtrip(tmp_path):
    root = h5.H5Group()
    root.attrs["title"] = "hello"
    root.attrs["ints"] = np.arange(5, dtype=np.int32)
    root.attrs["names"] = [b"a", b"bcd", b"ef"]
    g = root.require_group("grp/sub")
    g.children["x"] = h5.H5Dataset(np.arange(24, dtype=np.float32).reshape(2, 3, 4), {"unit": b"px"})
    g.children["y"] = h5.H5Dataset(np.arange(7, dtype=np.float64))
    for i in range(40):  # more links than one SNOD holds (2*leafK = 8)
        root.require_group("many").children[f"d{i:02d}"] = h5.H5Dataset(np.full((3,), i, np.int64))
    p = str(tmp_path / "t.h5")
    h5.write_h5(p, root)
    back = h5.read_h5(p)
    assert back.attrs["title"] == b"hello"
    np.testing.assert_array_equal(back.attrs["ints"], np.arange(5))
    assert [bytes(x) for x in back.attrs["names"]] == [b"a", b"bcd", b"ef"]
    np.testing.assert_array_equal(back["grp/sub/x"].data, root["grp/sub/x"].data)
    assert back["grp/sub/x"].attrs["unit"] == b"px"
    assert back["grp/sub/y"].data.dtype == np.float64
    assert len(back["many"].children) == 40 and int(back["many/d17"].data[0]) == 17


def test_file_structure_is_valid_hdf5_superblock_v0(tmp_path):
    p = str(tmp_path / "s.h5")
    h5.write_h5(p, h5.H5Group(children={"a": h5.H5Dataset(np.zeros(4, np.float32))}))
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8] == 0 and raw[13] == 8 and raw[14] == 8
    eof = struct.unpack_from("<Q", raw, 40)[0]
    assert eof == len(raw)
    assert b"TREE" in raw and b"HEAP" in raw and b"SNOD" in raw


def test_reader_handles_vlen_string_attribute():
    """h5py stores python-str attributes as variable-length strings in a global heap; hand-build one."""
    w = h5._Writer()
    s = b"tensorflow"
    gcol = b"GCOL" + bytes([1, 0, 0, 0]) + struct.pack("<Q", 16 + 16 + 16) + struct.pack("<HHIQ", 1, 0, 0, len(s)) + \
        s.ljust(16, b"\0")
    gaddr = w.alloc(gcol)
    dt = bytes([0x10 | 9, 0x01, 0x00, 0]) + struct.pack("<I", 16) + bytes([0x10 | 3, 0, 0, 0]) + struct.pack("<I", 1)
    ds = h5._dspace_msg(())
    nm = b"backend\0"
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + h5._pad8(nm) + h5._pad8(dt) + h5._pad8(ds) + \
        struct.pack("<IQI", len(s), gaddr, 1)
    heap_data = w.alloc(b"\0" * 8)
    heap = w.alloc(b"HEAP" + bytes(4) + struct.pack("<QQQ", 8, 1, heap_data))
    node = b"TREE" + struct.pack("<BBHQQ", 0, 0, 0, h5.UNDEF, h5.UNDEF) + b"\0" * (8 * 65)
    bt = w.alloc(node)
    oh = w.alloc(h5._object_header([h5._msg(0x11, struct.pack("<QQ", bt, heap)), h5._msg(0x0C, body)]))
    sb = h5.SIG + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", 4, 16, 0) + \
        struct.pack("<QQQQ", 0, h5.UNDEF, len(w.buf), h5.UNDEF) + struct.pack("<QQII", 0, oh, 1, 0) + struct.pack("<QQ", bt, heap)
    w.buf[0:96] = sb
    root = h5.read_h5_bytes(bytes(w.buf))
    assert root.attrs["backend"] == b"tensorflow"


@pytest.mark.parametrize("full_model", [True, False])
def test_keras_layout_roundtrip(tmp_path, full_model):
    w = S.init_weights(7)
    p = str(tmp_path / "crack_segmentation.h5")
    save_keras_h5(p, S, w, full_model=full_model)
    root = h5.read_h5(p)
    wroot = root["model_weights"] if full_model else root
    names = [bytes(x).decode() for x in wroot.attrs["layer_names"]]
    assert "separable_conv2d_5" in names and "conv2d_transpose_7" in names and "max_pooling2d_2" in names
    assert [bytes(x) for x in wroot["conv2d"].attrs["weight_names"]] == [b"conv2d/kernel:0", b"conv2d/bias:0"]
    assert wroot["separable_conv2d/separable_conv2d/depthwise_kernel:0"].data.shape == (3, 3, 32, 1)
    if full_model:
        cfg = __import__("json").loads(bytes(root.attrs["model_config"]).decode())
        assert cfg["class_name"] == "Functional" and len(cfg["config"]["layers"]) == len(keras_layer_sequence(S))
    back = load_keras_h5(p, S)
    for a, b in zip(w, back):
        np.testing.assert_array_equal(a, b)


def test_name_keyed_load_tolerates_session_suffix_offsets(tmp_path):
    """A model built second in a Keras session is named conv2d_9, batch_normalization_15, ...: still loads."""
    w = S.init_weights(8)
    root = h5.H5Group()
    off = {"conv2d": 9, "batch_normalization": 15, "separable_conv2d": 6, "conv2d_transpose": 8}
    names = []
    for l in S.layers:
        base = l.name.rsplit("_", 1)[0] if l.name.rsplit("_", 1)[-1].isdigit() else l.name
        idx = int(l.name.rsplit("_", 1)[1]) if l.name.rsplit("_", 1)[-1].isdigit() else 0
        new = f"{base}_{idx + off[base]}"
        names.append(new)
        g = root.require_group(new)
        wn = []
        for ti in l.tensors:
            t = S.tensors[ti]
            wn.append(f"{new}/{t.weight}:0".encode())
            g.require_group(new).children[f"{t.weight}:0"] = h5.H5Dataset(w[ti])
        g.attrs["weight_names"] = wn
    root.attrs["layer_names"] = [n.encode() for n in reversed(names)]  # order in the file must not matter
    p = str(tmp_path / "w.h5")
    h5.write_h5(p, root)
    back = load_keras_h5(p, S)
    for a, b in zip(w, back):
        np.testing.assert_array_equal(a, b)
    with pytest.raises((KeyError, ValueError)):
        del root.children[names[3]]
        h5.write_h5(p, root)
        load_keras_h5(p, S)
