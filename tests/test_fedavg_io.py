"""FedAvg math, weight formats, checkpoint round-trips."""
import pickle

import numpy as np
import pytest

from crackfl.io import checkpoint as ckpt
from crackfl.io import weights as wio
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.parallel import fedavg as fa


def _reference_updateWeight(received_parameters, round_client):
    """NumPy transcription of the arithmetic at fl_server.py:92-102 (single round)."""
    averaged = []
    for wcl in received_parameters:
        if len(averaged) == 0:
            averaged = [np.array(w) for w in wcl]
        else:
            for i, wc in enumerate(wcl):
                averaged[i] = averaged[i] + wc
    return [aw / round_client for aw in averaged]


def _rand_updates(n, seed=0):
    rng = np.random.default_rng(seed)
    return [[rng.standard_normal(t.shape).astype(np.float32) for t in S.tensors] for _ in range(n)]


def test_fedavg_mean_matches_reference_arithmetic():
    ups = _rand_updates(3)
    ours = fa.fedavg_mean(ups)
    ref = _reference_updateWeight(ups, 3)
    for a, b in zip(ours, ref):
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-6)


def test_fedavg_weighted_and_uniform_equivalence():
    ups = _rand_updates(4, 1)
    a = fa.fedavg_weighted(ups, [5, 5, 5, 5]); b = fa.fedavg_mean(ups)
    for x, y in zip(a, b):
        np.testing.assert_allclose(x, y, rtol=1e-5, atol=1e-6)
    c = fa.fedavg_weighted(ups[:2], [1, 3])
    np.testing.assert_allclose(c[0], 0.25 * ups[0][0] + 0.75 * ups[1][0], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(fa.normalise_coefficients(2, [1, 3]), [0.25, 0.75])
    with pytest.raises(ValueError):
        fa.fedavg_mean([])


def test_fedadam_step():
    ups = _rand_updates(2, 2)
    g = [np.zeros(t.shape, np.float32) for t in S.tensors]
    opt = fa.FedAdam(eta=0.1, trainable_mask=[t.trainable for t in S.tensors])
    out = opt.aggregate(g, ups, [1, 1])
    avg = fa.fedavg_mean(ups)
    i = 0
    d = avg[i]
    exp = 0.1 * (0.1 * d) / (np.sqrt(0.01 * d * d) + 1e-3)
    np.testing.assert_allclose(out[i], exp, rtol=1e-4, atol=1e-6)
    j = S.index("batch_normalization/moving_mean")
    np.testing.assert_allclose(out[j], avg[j])  # BN statistics are plain-averaged


def test_pickle_format_is_the_reference_wire_format():
    w = S.init_weights(0)
    blob = wio.dumps_pickle(w)
    plain = pickle.loads(blob)  # what the reference does (client_fit_model.py:51)
    assert isinstance(plain, list) and len(plain) == 112 and plain[0].dtype == np.float32
    back = wio.loads_pickle(pickle.dumps(plain))
    for a, b in zip(w, back):
        np.testing.assert_array_equal(a, b)


def test_restricted_unpickler_blocks_code_execution():
    class Evil:
        def __reduce__(self):
            import os
            return (os.system, ("echo pwned",))
    with pytest.raises(wio.UnsafePickleError):
        wio.loads_pickle(pickle.dumps(Evil()))
    with pytest.raises(wio.UnsafePickleError):
        wio.loads_pickle(pickle.dumps({"a": 1}))
    with pytest.raises(wio.UnsafePickleError):
        wio.loads_auto(wio.dumps_pickle(S.init_weights(0)), allow_pickle=False)


def test_raw_framing_roundtrip_and_truncation():
    w = S.init_weights(3)
    blob = wio.dumps_raw(w)
    assert len(blob) < S.num_params * 4 + 4096
    back = wio.loads_auto(blob)
    for a, b in zip(w, back):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ValueError):
        wio.loads_raw(blob[:-10])


def test_flatten_unflatten_and_checkpoint(tmp_path):
    w = S.init_weights(4)
    flat = wio.flatten(w, S)
    assert flat.size == S.arena_numel and flat.size % 1024 == 0
    for a, b in zip(w, wio.unflatten(flat, S)):
        np.testing.assert_array_equal(a, b)
    m = [x * 0 + 1 for x in w]
    ckpt.save_checkpoint(str(tmp_path), S, w, meta={"current_round": 3}, extra_arenas={"fedadam_m": m})
    st = ckpt.load_checkpoint(str(tmp_path), S)
    assert st["meta"]["current_round"] == 3
    np.testing.assert_array_equal(st["weights"][5], w[5])
    np.testing.assert_array_equal(st["arenas"]["fedadam_m"][7], m[7])
    with pytest.raises(ValueError):
        wio.flatten(w[:-1], S)


def test_checkpoint_commits_at_the_manifest(tmp_path, monkeypatch):
    """Arena files carry a save id and the manifest rename is the single commit point: a crash after the new arenas are
    written but before the manifest is replaced must leave the OLD round with the OLD weights."""
    from crackfl.io import checkpoint as ck
    from crackfl.models.spec import REFERENCE_SPEC as S
    w1 = S.init_weights(1)
    w2 = [w + 1 for w in w1]
    d = str(tmp_path / "ck")
    ck.save_checkpoint(d, S, w1, meta={"current_round": 2, "model_version": 2})
    real = ck._atomic_write
    def crash_on_manifest(path, data):
        if path.endswith("manifest.json"):
            raise OSError("power cut")
        real(path, data)
    monkeypatch.setattr(ck, "_atomic_write", crash_on_manifest)
    with pytest.raises(OSError):
        ck.save_checkpoint(d, S, w2, meta={"current_round": 3, "model_version": 3})
    monkeypatch.setattr(ck, "_atomic_write", real)
    st = ck.load_checkpoint(d, S)
    assert st["meta"]["current_round"] == 2
    np.testing.assert_array_equal(st["weights"][0], w1[0])
    # a completed save replaces it and removes the previous arenas
    ck.save_checkpoint(d, S, w2, meta={"current_round": 3, "model_version": 3})
    st = ck.load_checkpoint(d, S)
    assert st["meta"]["current_round"] == 3
    np.testing.assert_array_equal(st["weights"][0], w2[0])
    import os
    assert sum(f.startswith("global.") for f in os.listdir(d)) <= 2  # the live one (+ the orphan of the crashed save)
