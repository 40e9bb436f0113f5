"""Property tests (hypothesis) of the pieces every round passes through: aggregation coefficients, the weighted mean and the
raw tensor framing.  Small random tensor lists instead of the 112-tensor model, so hundreds of cases run in seconds."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from crackfl.io import weights as wio
from crackfl.parallel import fedavg as fa

shapes = st.lists(st.lists(st.integers(1, 5), min_size=0, max_size=3).map(tuple), min_size=1, max_size=5)


def _updates(shape_list, k, seed):
    rng = np.random.default_rng(seed)
    return [[rng.standard_normal(s).astype(np.float32) for s in shape_list] for _ in range(k)]


@settings(max_examples=60, deadline=None)
@given(shapes, st.integers(1, 6), st.integers(0, 2 ** 16))
def test_uniform_mean_is_the_reference_sum_over_count(shape_list, k, seed):
    """fl_server.py:92-102: sum the clients' lists element-wise, divide by the number of clients."""
    ups = _updates(shape_list, k, seed)
    got = fa.fedavg_mean(ups)
    for i, s in enumerate(shape_list):
        want = sum(u[i].astype(np.float64) for u in ups) / k
        assert got[i].shape == tuple(s) and got[i].dtype == np.float32
        np.testing.assert_allclose(got[i], want, rtol=1e-5, atol=1e-6)


@settings(max_examples=60, deadline=None)
@given(shapes, st.lists(st.integers(1, 10 ** 6), min_size=1, max_size=6), st.integers(0, 2 ** 16))
def test_weighted_mean_properties(shape_list, ns, seed):
    k = len(ns)
    ups = _updates(shape_list, k, seed)
    c = fa.normalise_coefficients(k, ns)
    assert abs(float(c.sum()) - 1.0) < 1e-6 and (c > 0).all()
    got = fa.fedavg_weighted(ups, ns)
    for i in range(len(shape_list)):
        want = sum(ci * u[i].astype(np.float64) for ci, u in zip(c, ups))
        np.testing.assert_allclose(got[i], want, rtol=1e-4, atol=1e-5)
        # a convex combination stays inside the clients' range
        lo = np.min([u[i] for u in ups], axis=0); hi = np.max([u[i] for u in ups], axis=0)
        assert (got[i] >= lo - 1e-5).all() and (got[i] <= hi + 1e-5).all()
    # scaling every sample count by the same factor changes nothing
    again = fa.fedavg_weighted(ups, [7 * n for n in ns])
    for a, b in zip(got, again):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)
    # identical clients: the mean is that client, whatever the counts
    same = fa.fedavg_weighted([ups[0]] * k, ns)
    for a, b in zip(same, ups[0]):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)


@settings(max_examples=60, deadline=None)
@given(shapes, st.integers(0, 2 ** 16))
def test_raw_framing_roundtrip_and_rejects_truncation(shape_list, seed):
    w = _updates(shape_list, 1, seed)[0]
    buf = wio.dumps_raw(w)
    back = wio.loads_raw(buf)
    assert len(back) == len(w)
    for a, b in zip(w, back):
        assert a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a, b)
    assert wio.loads_auto(buf) is not None
    if len(buf) > 1:
        with pytest.raises(Exception):
            wio.loads_raw(buf[:-1])


# ---- pure-Python HDF5 subset: random trees survive a write -> read round trip
from crackfl.io import hdf5 as h5  # noqa: E402

_names = st.text(alphabet="abcdefghijklmnopqrstuvwxyzABCXYZ0123456789_:.", min_size=1, max_size=24)
_dtypes = st.sampled_from([np.float32, np.float64, np.int32, np.int64, np.uint8])
_arrays = st.tuples(_dtypes, st.lists(st.integers(0, 4), min_size=0, max_size=4), st.integers(0, 2 ** 16))


def _mk(dt, shape, seed):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(shape) * 50).astype(dt)


@settings(max_examples=60, deadline=None)
@given(st.dictionaries(_names, st.dictionaries(_names, _arrays, min_size=0, max_size=12), min_size=0, max_size=4),
       st.dictionaries(_names, _arrays, min_size=0, max_size=3))
def test_hdf5_random_tree_roundtrip(groups, top):
    root = h5.H5Group()
    root.attrs["n_groups"] = np.array([len(groups)], np.int64)
    for gname, dsets in groups.items():
        g = root.require_group(gname)
        g.attrs["label"] = gname.encode()
        for dname, (dt, shape, seed) in dsets.items():
            g.children[dname] = h5.H5Dataset(_mk(dt, shape, seed), {"seed": np.array([seed], np.int32)})
    for dname, (dt, shape, seed) in top.items():
        if dname not in root.children:
            root.children[dname] = h5.H5Dataset(_mk(dt, shape, seed))
    w = h5._Writer()
    back = h5.read_h5_bytes(w.finish(root))
    assert int(back.attrs["n_groups"][0]) == len(groups)
    for gname, dsets in groups.items():
        g = back[gname]
        assert bytes(g.attrs["label"]) == gname.encode()
        assert set(g.children) >= set(dsets)
        for dname, (dt, shape, seed) in dsets.items():
            want = _mk(dt, shape, seed)
            got = g.children[dname].data
            assert got.dtype == want.dtype and got.shape == want.shape and np.array_equal(got, want)
            assert int(g.children[dname].attrs["seed"][0]) == seed
    for dname, (dt, shape, seed) in top.items():
        if isinstance(back.children.get(dname), h5.H5Dataset):
            want = _mk(dt, shape, seed)
            got = back.children[dname].data
            assert got.dtype == want.dtype and got.shape == want.shape and np.array_equal(got, want)


# ---- server round state machine under random interleavings (fl_server.py:107-150 semantics + the documented fixes)
from crackfl.config import fast_test_config  # noqa: E402
from crackfl.models.spec import REFERENCE_SPEC as S  # noqa: E402
from crackfl.protocol import client as C  # noqa: E402
from crackfl.protocol import messages as pb  # noqa: E402
from crackfl.protocol.server import FLServer  # noqa: E402

_W0 = S.init_weights(0)


class _Clock:
    def __init__(self):
        self.t = 0.0

    def __call__(self):
        return self.t

    def sleep(self, s):
        self.t += s


@settings(max_examples=12, deadline=None)
@given(st.integers(1, 4), st.integers(1, 3), st.randoms(use_true_random=False))
def test_round_state_machine_under_random_interleavings(tmp_path_factory, k, rounds, rnd):
    tmp = tmp_path_factory.mktemp("sm")
    cfg = fast_test_config()
    cfg.server.weights_dir = str(tmp / "w"); cfg.server.log_dir = str(tmp / "l"); cfg.server.max_rounds = rounds
    clk = _Clock()
    srv = FLServer(cfg, [w.copy() for w in _W0], S, clock=clk, sleep=clk.sleep, log=lambda s: None)

    def one(req):
        out = srv.handle(req)
        assert len(out) == 1
        return out[0]

    names = []
    for i in range(k):
        c = pb.from_config(one(next(C.request_ready(f"c{i}"))).ready_rep.config)
        assert c["state"] == "SW"
        names.append(c["cname"])
    finished = set()
    for r in range(1, rounds + 1):
        order = names[:]
        rnd.shuffle(order)
        offs = {n: float(rnd.randint(-3, 3)) for n in names}
        want0 = _W0[0] + np.float32(sum(offs.values()) / k)
        for pos, n in enumerate(order):
            if r > 1 and rnd.random() < 0.3:  # a stale upload from the previous round changes nothing
                reqs = list(C.request_traindone(n, r - 1, wio.dumps_raw(_W0), {}, 1 << 30))
                assert pb.from_config(one(reqs[0]).update_rep.config)["state"] == "RESP_ACY"
            if rnd.random() < 0.5:  # a version poll in the middle of the round: told to wait
                v = one(next(C.request_model_version(r, r))).version_rep
                assert v.state == pb.State.WAIT
            w = [x + np.float32(offs[n]) if i == 0 else x for i, x in enumerate(_W0)]
            reqs = list(C.request_traindone(n, r, wio.dumps_raw(w), {}, 1 << 30))
            rep = one(reqs[0]).update_rep
            state = pb.from_config(rep.config)["state"]
            last = pos == k - 1
            if not last:
                assert state == "RESP_ACY" and rep.buffer_chunk == b""
            elif r < rounds:
                assert state == "RESP_ARY"
                np.testing.assert_allclose(wio.loads_auto(rep.buffer_chunk)[0], want0, rtol=1e-5, atol=1e-5)
            else:
                assert state == "FIN"
                finished.add(n)
        np.testing.assert_allclose(srv.global_weights[0], want0, rtol=1e-5, atol=1e-5)
        assert srv.model_version == cfg.server.initial_model_version + r and len(srv.history) == r
        for n in order[:-1]:  # the clients that were told to wait learn the outcome from the poll
            v = one(next(C.request_model_version(r, r))).version_rep
            if r < rounds:
                assert v.state == pb.State.NOT_WAIT and pb.from_config(v.config)["current_round"] == r + 1
                np.testing.assert_allclose(wio.loads_auto(v.buffer_chunk)[0], want0, rtol=1e-5, atol=1e-5)
            else:
                assert v.state == pb.State.FIN
                finished.add(n)
    assert finished == set(names) and srv.finished


@settings(max_examples=80, deadline=None)
@given(st.binary(min_size=0, max_size=4000), st.integers(1, 700), st.randoms(use_true_random=False))
def test_chunked_upload_reassembles_in_any_order(tmp_path_factory, payload, cap, rnd):
    tmp = tmp_path_factory.mktemp("ch")
    cfg = fast_test_config()
    cfg.server.weights_dir = str(tmp / "w"); cfg.server.log_dir = str(tmp / "l")
    srv = FLServer(cfg, [w.copy() for w in _W0], S, log=lambda s: None)
    reqs = list(C.request_traindone("c0", 1, payload, {"num_samples": 3}, cap))
    assert b"".join(r.update_req.buffer_chunk for r in reqs) == payload
    order = list(range(len(reqs)))
    rnd.shuffle(order)
    if len(order) > 1 and rnd.random() < 0.5:
        order.insert(rnd.randrange(len(order)), order[0])  # one chunk arrives twice
    out = None
    for pos, i in enumerate(order):
        u = reqs[i].update_req
        out = srv._assemble_chunks(u.cname, u.buffer_chunk, pb.from_config(u.config))
        if pos < len(order) - 1 and len(reqs) > 1:
            assert out is None or pos >= len(set(order[:pos + 1])) - 1
    assert out == payload
    assert "c0" not in srv._chunks  # nothing is left behind


def test_chunk_headers_that_cannot_be_right_are_refused(tmp_path):
    cfg = fast_test_config()
    cfg.server.weights_dir = str(tmp_path / "w"); cfg.server.log_dir = str(tmp_path / "l")
    srv = FLServer(cfg, [w.copy() for w in _W0], S, log=lambda s: None)
    for hdr in ({"chunk_index": 3, "chunk_count": 3}, {"chunk_index": -1, "chunk_count": 2}, {"chunk_index": 0, "chunk_count": 10 ** 6}):
        with pytest.raises(ValueError):
            srv._assemble_chunks("c0", b"x", hdr)
    # a half-sent upload is superseded by a fresh one with another chunk count, and by an unchunked payload
    assert srv._assemble_chunks("c0", b"a", {"chunk_index": 0, "chunk_count": 3}) is None
    assert srv._assemble_chunks("c0", b"b", {"chunk_index": 0, "chunk_count": 2}) is None
    assert srv._assemble_chunks("c0", b"c", {"chunk_index": 1, "chunk_count": 2}) == b"bc"
    assert srv._assemble_chunks("c0", b"a", {"chunk_index": 0, "chunk_count": 3}) is None
    assert srv._assemble_chunks("c0", b"whole", {}) == b"whole" and "c0" not in srv._chunks


@settings(max_examples=200, deadline=None)
@given(st.dictionaries(st.text(min_size=1, max_size=12),
                       st.one_of(st.booleans(), st.integers(-2 ** 40, 2 ** 40), st.floats(allow_nan=False, width=32),
                                 st.text(max_size=20),
                                 st.integers(-1000, 1000).map(np.int64), st.integers(0, 255).map(np.uint8),
                                 st.floats(-1e3, 1e3).map(np.float32), st.booleans().map(np.bool_)), max_size=8))
def test_config_map_codec_never_raises_and_keeps_values(d):
    """The Scalar map of every message (fl_server.py:61-73): whatever a trainer puts into its metrics must get across."""
    back = pb.from_config(pb.to_config(d))
    assert set(back) == set(d)
    for k, v in d.items():
        b = back[k]
        if isinstance(v, (bool, np.bool_)):
            assert b == int(v)
        elif isinstance(v, (int, np.integer)):
            assert int(b) == int(v)  # out-of-int32 values travel as their decimal string
        elif isinstance(v, (float, np.floating)):
            assert isinstance(b, float) and (b == np.float32(v) or abs(b - float(v)) <= 1e-6 * abs(float(v)))
        else:
            assert b == v


def test_config_json_round_trip_and_override_types(tmp_path):
    """Every knob survives to_dict -> JSON -> load, and a --section.key=value override keeps the field's type."""
    import dataclasses
    import json as _json
    from crackfl.config import Config
    base = Config()
    base.train.image_size = (96, 160); base.server.round_timeout_s = 12.5; base.train.validate = False
    p = tmp_path / "c.json"
    p.write_text(_json.dumps(base.to_dict()))
    back = Config.load(str(p), env=False)
    assert back.to_dict() == base.to_dict() and back.train.image_size == (96, 160)
    for sec_f in dataclasses.fields(base):
        sec = getattr(base, sec_f.name)
        for f in dataclasses.fields(sec):
            cur = getattr(sec, f.name)
            if isinstance(cur, bool):
                txt, want = ("false" if cur else "true"), (not cur)
            elif isinstance(cur, int):
                txt, want = str(cur + 3), cur + 3
            elif isinstance(cur, float):
                txt, want = repr(cur + 0.25), cur + 0.25
            elif isinstance(cur, tuple):
                txt, want = "48x80", (48, 80)
            else:
                txt, want = "some/value", "some/value"
            c = Config.load(None, overrides=[f"--{sec_f.name}.{f.name}={txt}"], env=False)
            got = getattr(getattr(c, sec_f.name), f.name)
            assert got == want and type(got) is type(want), (sec_f.name, f.name, got)
    with pytest.raises(KeyError):
        Config.load(None, overrides=["--train.no_such_knob=1"], env=False)
    with pytest.raises(KeyError):
        Config.load(None, overrides=["--nosection.x=1"], env=False)


@settings(max_examples=80, deadline=None)
@given(st.integers(0, 60), st.integers(0, 70), st.integers(0, 2 ** 16), st.sampled_from(["client", "script"]))
def test_split_keeps_pairs_together_and_partitions_the_set(n, count, seed, mode):
    """client_fit_model.py:76-82 / test/Segmentation.py:84-90: two same-seed shuffles, then a cut."""
    from crackfl.data.generator import split_paths
    imgs = [f"img/{i:04d}.jpg" for i in range(n)]
    msks = [f"msk/{i:04d}.jpg" for i in range(n)]
    tx, ty, vx, vy = split_paths(imgs, msks, count, seed, mode)
    assert len(tx) == len(ty) and len(vx) == len(vy)
    for a, b in zip(tx + vx, ty + vy):
        assert a[4:] == b[4:]  # image i stays with mask i
    if count > 0 or mode == "client":
        assert sorted(tx + vx) == imgs and not set(tx) & set(vx)
    if mode == "client":
        assert len(tx) == min(count, n)
    elif count > 0:
        assert len(vx) == min(count, n)
    with pytest.raises(ValueError):
        split_paths(imgs + ["img/extra.jpg"], msks, count, seed, mode)
