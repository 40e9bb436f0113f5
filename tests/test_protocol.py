"""Protocol state machine + gRPC loopback (BASELINE config 1)."""
import os
import re
import threading

import numpy as np
import pytest

from crackfl.config import Config, fast_test_config
from crackfl.io import weights as wio
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.protocol import messages as pb
from crackfl.protocol import client as C
from crackfl.protocol.server import FLServer, FaultInjector, make_grpc_server


class FakeClock:
    def __init__(self):
        self.t = 0.0
    def __call__(self):
        return self.t
    def sleep(self, s):
        self.t += s


def _server(tmp_path, **kw):
    cfg = fast_test_config()
    cfg.server.weights_dir = str(tmp_path / "server_weights")
    cfg.server.log_dir = str(tmp_path / "logs")
    for k, v in kw.items():
        setattr(cfg.server, k, v)
    clk = FakeClock()
    srv = FLServer(cfg, S.init_weights(0), S, clock=clk, sleep=clk.sleep, log=lambda s: None)
    return srv, clk, cfg


def _one(srv, req):
    out = srv.handle(req)
    assert len(out) == 1
    return out[0]


def _ready(srv, name):
    return pb.from_config(_one(srv, next(C.request_ready(name))).ready_rep.config)


def _done(srv, name, cr, w, extra=None):
    reqs = list(C.request_traindone(name, cr, wio.dumps_raw(w), extra or {}, 1 << 30))
    rep = _one(srv, reqs[0]).update_rep
    return pb.from_config(rep.config), rep.buffer_chunk


def test_proto_text_matches_runtime_descriptors():
    txt = open(os.path.join(os.path.dirname(pb.__file__), "transport.proto")).read()
    fd = pb.FILE_DESCRIPTOR_PROTO
    for m in fd.message_type:
        assert re.search(rf"message {m.name}\s*{{", txt), m.name
        for f in m.field:
            assert re.search(rf"\b{f.name}\s*=\s*{f.number};", txt), (m.name, f.name)
    for v in fd.enum_type[0].value:
        assert re.search(rf"\b{v.name}\s*=\s*{v.number};", txt)
    assert "rpc transport(stream transportRequest) returns (stream transportResponse)" in txt
    assert pb.VersionRep(state="FIN").state == pb.State.FIN  # fl_server.py:145 passes the string


def test_admission_window_sw_then_ctw(tmp_path):
    srv, clk, cfg = _server(tmp_path, admission_window_s=10.0, ready_delay_s=5.0)
    c = _ready(srv, "a")
    assert c["state"] == "SW" and c["current_round"] == 1 and c["max_train_round"] == 5 and c["model_version"] == 1
    assert clk.t == 5.0  # the per-ready delay of fl_server.py:56
    assert _ready(srv, "b")["state"] == "SW"   # t=5 -> 10: still inside the 10 s window at entry
    assert _ready(srv, "late")["state"] == "CTW"  # t=10: window closed
    assert srv.ready_clients == ["a", "b"]


def test_name_collision_gets_unique_id(tmp_path):
    srv, _, _ = _server(tmp_path)
    a = _ready(srv, "client7"); b = _ready(srv, "client7")
    assert a["cname"] == "client7" and b["cname"] != "client7" and len(srv.ready_clients) == 2


def test_round_flow_acy_ary_fin_and_version_poll(tmp_path):
    srv, clk, cfg = _server(tmp_path, max_rounds=2)
    _ready(srv, "a"); _ready(srv, "b")
    rep = _one(srv, next(C.request_parameter())).update_rep
    assert rep.type == "P" and rep.title == "parameters"
    w0 = wio.loads_auto(rep.buffer_chunk)
    assert _one(srv, next(C.request_training("a"))).update_rep.type == "T"
    wa = [x + 1 for x in w0]; wb = [x + 3 for x in w0]
    c, blob = _done(srv, "a", 1, wa)
    assert c["state"] == "RESP_ACY" and blob == b""
    v = _one(srv, next(C.request_model_version(1, 1))).version_rep
    assert v.state == pb.State.WAIT
    c, blob = _done(srv, "b", 1, wb)
    assert c["state"] == "RESP_ARY" and c["current_round"] == 2 and c["model_version"] == 2
    agg = wio.loads_auto(blob)
    np.testing.assert_allclose(agg[0], w0[0] + 2, rtol=1e-6)  # mean of the two updates
    v = _one(srv, next(C.request_model_version(1, 1))).version_rep
    assert v.state == pb.State.NOT_WAIT and pb.from_config(v.config)["current_round"] == 2
    np.testing.assert_allclose(wio.loads_auto(v.buffer_chunk)[3], agg[3])
    # stale round is dropped with RESP_ACY (fl_server.py:134-135)
    c, _ = _done(srv, "a", 1, wa)
    assert c["state"] == "RESP_ACY" and len(srv.round_updates) == 0
    # round 2 = last round -> FIN; the straggler learns FIN from the poll (fix Q7)
    c, _ = _done(srv, "a", 2, wa); assert c["state"] == "RESP_ACY"
    c, _ = _done(srv, "b", 2, wb); assert c["state"] == "FIN"
    v = _one(srv, next(C.request_model_version(2, 2))).version_rep
    assert v.state == pb.State.FIN
    # per-round mean only (fix Q1): round 2 result is again the mean of exactly two updates
    np.testing.assert_allclose(srv.global_weights[0], w0[0] + 2, rtol=1e-6)
    assert os.path.exists(os.path.join(cfg.server.weights_dir, "weights.pickle"))
    assert len(srv.history) == 2


def test_sample_weighted_strategy(tmp_path):
    srv, _, _ = _server(tmp_path, strategy="fedavg_weighted")
    _ready(srv, "a"); _ready(srv, "b")
    w0 = S.init_weights(0)
    _done(srv, "a", 1, [x * 0 + 1 for x in w0], {"num_samples": 1})
    c, blob = _done(srv, "b", 1, [x * 0 + 5 for x in w0], {"num_samples": 3})
    assert c["state"] == "RESP_ARY"
    np.testing.assert_allclose(wio.loads_auto(blob)[0], 4.0, rtol=1e-6)


def test_round_deadline_with_quorum(tmp_path):
    srv, clk, _ = _server(tmp_path, round_timeout_s=30.0, min_fit_clients=1)
    _ready(srv, "a"); _ready(srv, "dead")
    w0 = S.init_weights(0)
    c, _ = _done(srv, "a", 1, w0)
    assert c["state"] == "RESP_ACY"
    clk.t += 31
    v = _one(srv, next(C.request_model_version(1, 1))).version_rep
    assert v.state == pb.State.NOT_WAIT and srv.ready_clients == ["a"]


def test_fault_injection_drop_and_duplicate(tmp_path):
    srv, _, _ = _server(tmp_path)
    srv.fault = FaultInjector(drop={"D": 1}, duplicate={"D": 1})
    _ready(srv, "a"); _ready(srv, "b")
    w0 = S.init_weights(0)
    req = list(C.request_traindone("a", 1, wio.dumps_raw(w0), {}, 1 << 30))[0]
    assert srv.handle(req) == []            # dropped
    c = pb.from_config(_one(srv, req).update_rep.config)  # duplicated delivery counts once
    assert c["state"] == "RESP_ACY" and len(srv.round_updates) == 1


def test_log_upload_is_confined(tmp_path):
    srv, _, cfg = _server(tmp_path)
    f = tmp_path / "ev.bin"; f.write_bytes(b"x" * 1000)
    for req in C.get_file_chunks(str(f), 400):
        req.update_req.title = "./send_logs/logs/r1/ev.bin"
        assert _one(srv, req).update_rep.type == "L"
    assert (tmp_path / "logs" / "logs" / "r1" / "ev.bin").read_bytes() == b"x" * 1000
    p = srv.save_log_chunk(b"z", "../../etc/passwd", False)
    assert p.startswith(os.path.abspath(cfg.server.log_dir))


def test_chunked_weights_both_directions(tmp_path):
    srv, _, cfg = _server(tmp_path)
    cfg.transport.max_message_mib = 3  # forces ~8.2 MB payloads into chunks
    _ready(srv, "a")
    reps = srv.handle(next(C.request_parameter()))
    assert len(reps) > 1
    _, blob = C._collect_weights(reps, "update_rep")
    w = wio.loads_auto(blob)
    reqs = list(C.request_traindone("a", 1, wio.dumps_raw(w), {}, 2 << 20))
    assert len(reqs) > 1
    states = [pb.from_config(_one(srv, r).update_rep.config)["state"] if i < len(reqs) - 1 else None
              for i, r in enumerate(reqs[:-1])]
    assert all(s == "CHUNK" for s in states)
    last = srv.handle(reqs[-1])
    assert pb.from_config(last[0].update_rep.config)["state"] == "RESP_ARY"


def test_server_resume(tmp_path):
    srv, _, cfg = _server(tmp_path)
    _ready(srv, "a")
    _done(srv, "a", 1, [x + 1 for x in S.init_weights(0)])
    cfg.server.resume = True
    srv2 = FLServer(cfg, S.init_weights(5), S, log=lambda s: None)
    assert srv2.current_round == 2 and srv2.model_version == 2
    np.testing.assert_allclose(srv2.global_weights[0], srv.global_weights[0])


class _StubTrainer:
    """Stands in for the local fit: adds a client-specific constant to every tensor."""
    def __init__(self, delta):
        self.delta, self.calls = delta, []
    def fit(self, weights, cr):
        self.calls.append(cr)
        return [w + self.delta for w in weights], {"loss": 0.5, "accuracy": 0.9, "num_samples": 16}


def test_two_clients_over_real_grpc_loopback(tmp_path, monkeypatch):
    """BASELINE config 1 plumbing: 2 clients, real gRPC on an ephemeral port, 2 rounds."""
    monkeypatch.chdir(tmp_path)
    cfg = fast_test_config()
    cfg.server.max_rounds = 2
    cfg.server.weights_dir = str(tmp_path / "sw")
    srv = FLServer(cfg, S.init_weights(0), S, log=lambda s: None)
    gs, port = make_grpc_server(srv, "127.0.0.1:0")
    gs.start()
    try:
        results = {}
        def run_client(i, delta):
            ccfg = fast_test_config()
            ccfg.client.name = f"c{i}"
            ccfg.client.weights_path = str(tmp_path / f"saved_weight_{i}" / "weights.pickle")
            call, ch = C.make_stub(ccfg, f"127.0.0.1:{port}")
            cl = C.FLClient(ccfg, _StubTrainer(delta), call, log=lambda s: None)
            results[i] = (cl.run(), cl)
            ch.close()
        ts = [threading.Thread(target=run_client, args=(i, d)) for i, d in ((0, 1.0), (1, 3.0))]
        [t.start() for t in ts]; [t.join(60) for t in ts]
        assert all(not t.is_alive() for t in ts)
    finally:
        gs.stop(0)
    assert results[0][0]["rounds"] == 2 and results[1][0]["rounds"] == 2
    w0 = S.init_weights(0)
    # each round adds mean(1,3)=2 to every tensor
    np.testing.assert_allclose(srv.global_weights[0], w0[0] + 4, rtol=1e-5)
    np.testing.assert_allclose(results[0][1].final_weights[0], w0[0] + 4, rtol=1e-5)
    assert srv.finished and len(srv.history) == 2
    assert srv.history[0]["client_metrics"][0]["accuracy"] == pytest.approx(0.9)
    assert os.path.exists(tmp_path / "saved_weight_0" / "weights.pickle")


def test_late_client_gets_ctw(tmp_path):
    srv, clk, cfg = _server(tmp_path)
    _ready(srv, "a")
    clk.t += 100
    cl = C.FLClient(cfg, _StubTrainer(0.0), srv.transport, sleep=clk.sleep, log=lambda s: None)
    assert cl.run()["state"] == "CTW"


def test_join_after_first_aggregation_does_not_reset_round(tmp_path):
    """A fast first round can finish inside the admission window.  A client admitted after it must get the LIVE round
    and model version (the reference would set the round back to 1, fl_server.py:61-63, and strand everybody)."""
    srv, clk, cfg = _server(tmp_path, admission_window_s=10.0, ready_delay_s=0.0, max_rounds=2)
    w0 = S.init_weights(0)
    assert _ready(srv, "a")["current_round"] == 1
    c, _ = _done(srv, "a", 1, [x + 1 for x in w0])
    assert c["state"] == "RESP_ARY" and srv.current_round == 2 and srv.model_version == 2
    clk.t = 3.0  # still inside the window
    b = _ready(srv, "b")
    assert b["state"] == "SW" and b["current_round"] == 2 and b["model_version"] == 2
    assert srv.current_round == 2  # not reset
    # both finish round 2: a is first (waits), b completes the round -> FIN for everybody
    c, _ = _done(srv, "a", 2, [x + 1 for x in w0]); assert c["state"] == "RESP_ACY"
    c, _ = _done(srv, "b", 2, [x + 3 for x in w0]); assert c["state"] == "FIN"
    v = _one(srv, next(C.request_model_version(2, 2, "a"))).version_rep
    assert v.state == pb.State.FIN and srv.all_fin_delivered()


def test_malformed_update_is_rejected_and_does_not_wedge_the_round(tmp_path):
    srv, clk, cfg = _server(tmp_path, max_rounds=1)
    _ready(srv, "good"); _ready(srv, "bad")
    w0 = S.init_weights(0)
    # broadcastable shape: NumPy would silently accept it and corrupt the mean
    wbad = [x.copy() for x in w0]; wbad[0] = np.zeros((1,), np.float32)
    with pytest.raises(ValueError, match="shape"):
        srv.manage_rounds("bad", 1, wbad)
    wint = [x.copy() for x in w0]; wint[1] = wint[1].astype(np.int32)
    with pytest.raises(ValueError, match="dtype"):
        srv.manage_rounds("bad", 1, wint)
    wnan = [x.copy() for x in w0]; wnan[2] = wnan[2] + np.nan
    with pytest.raises(ValueError, match="non-finite"):
        srv.manage_rounds("bad", 1, wnan)
    assert srv.round_updates == [] and srv.round_reporters == []
    # the round still completes once proper updates arrive
    c, _ = _done(srv, "good", 1, [x + 1 for x in w0]); assert c["state"] == "RESP_ACY"
    c, _ = _done(srv, "bad", 1, [x + 3 for x in w0]); assert c["state"] == "FIN"
    np.testing.assert_allclose(srv.global_weights[0], w0[0] + 2, rtol=1e-6)


def test_aggregator_failure_evicts_the_triggering_update(tmp_path):
    srv, clk, cfg = _server(tmp_path, max_rounds=1)

    class Flaky:
        kind, calls = "host", 0
        def aggregate(self, g, ups, ns=None):
            Flaky.calls += 1
            if Flaky.calls == 1:
                raise RuntimeError("boom")
            return [np.mean([u[i] for u in ups], axis=0) for i in range(len(g))]
    srv.aggregator = Flaky()
    _ready(srv, "a"); _ready(srv, "b")
    w0 = S.init_weights(0)
    _done(srv, "a", 1, [x + 1 for x in w0])
    with pytest.raises(RuntimeError):
        srv.manage_rounds("b", 1, [x + 3 for x in w0])
    assert srv.round_reporters == ["a"] and "b" not in srv.ready_clients  # evicted, not queued
    # nothing is wedged: a is the only client waited for; its (already queued) update aggregates on the next trigger
    c, _ = _done(srv, "c-late", 1, [x + 5 for x in w0])
    assert c["state"] == "FIN" and srv.finished


def test_admission_capped_at_aggregator_slots(tmp_path):
    srv, clk, cfg = _server(tmp_path)
    srv.aggregator.max_clients = 2
    assert _ready(srv, "a")["state"] == "SW" and _ready(srv, "b")["state"] == "SW"
    assert _ready(srv, "c")["state"] == "CTW" and len(srv.ready_clients) == 2


def test_serve_forever_keeps_serving_until_every_client_heard_fin(tmp_path, monkeypatch):
    """serve_forever used to stop 2 s after the last aggregation; clients still inside their poll interval then hit a
    dead server.  Here the poll interval (1.5 s) is longer than that old drain would be relative to the test's clock."""
    import socket
    import time as _t
    from crackfl.protocol.server import serve_forever
    monkeypatch.chdir(tmp_path)
    s_ = socket.socket(); s_.bind(("127.0.0.1", 0)); port = s_.getsockname()[1]; s_.close()
    cfg = fast_test_config()
    cfg.server.max_rounds = 1
    cfg.server.weights_dir = str(tmp_path / "sw")
    cfg.transport.bind = f"127.0.0.1:{port}"
    cfg.client.poll_interval_s = 1.5
    cfg.server.fin_grace_s = 20.0
    srv = FLServer(cfg, S.init_weights(0), S, log=lambda s: None)
    st = threading.Thread(target=serve_forever, args=(srv,), daemon=True)
    st.start()
    _t.sleep(0.3)
    results = {}
    def run_client(i, delay):
        ccfg = fast_test_config()
        ccfg.client.name = f"c{i}"
        ccfg.client.poll_interval_s = 1.5
        ccfg.client.write_weights_file = False
        call, ch = C.make_stub(ccfg, f"127.0.0.1:{port}")
        tr = _StubTrainer(1.0)
        orig = tr.fit
        def slow_fit(w, cr):
            _t.sleep(delay)
            return orig(w, cr)
        tr.fit = slow_fit
        cl = C.FLClient(ccfg, tr, call, log=lambda s: None)
        results[i] = (cl.run(), cl.final_weights is not None)
        ch.close()
    ts = [threading.Thread(target=run_client, args=(i, d)) for i, d in ((0, 0.0), (1, 0.4))]
    [t.start() for t in ts]; [t.join(60) for t in ts]
    st.join(30)
    assert not st.is_alive()
    assert results[0][0]["state"] == "FIN" and results[1][0]["state"] == "FIN"
    assert results[0][1], "the first reporter must still be able to fetch the final model"
    assert srv.all_fin_delivered()


def test_fedadam_on_the_host_aggregator_runs_past_round_one_and_survives_a_failed_aggregation(tmp_path):
    """The guarded aggregation snapshots the server optimiser's moments before every round: per-tensor lists of different
    shapes (a ragged np.array() of them raised from round 2 on), restored when the aggregation fails."""
    from crackfl.parallel import fedavg as fa
    srv, _, cfg = _server(tmp_path, strategy="fedadam", max_rounds=4, aggregator="host")
    _ready(srv, "a")
    w0 = S.init_weights(0)
    ref = fa.FedAdam(eta=cfg.server.fedadam_eta, beta1=cfg.server.fedadam_beta1, beta2=cfg.server.fedadam_beta2,
                     tau=cfg.server.fedadam_tau, trainable_mask=[t.trainable for t in S.tensors])
    g = [w.copy() for w in w0]
    for r in (1, 2):
        up = [x + np.float32(0.01 * r) for x in w0]
        c, _ = _done(srv, "a", r, up, {"num_samples": 4})
        assert c["state"] == "RESP_ARY" and c["current_round"] == r + 1
        g = ref.aggregate(g, [up], [4])
        np.testing.assert_allclose(srv.global_weights[0], g[0], rtol=1e-5, atol=1e-6)
    # round 3: the aggregator fails once; the moments must be what they were before the attempt
    strat = srv.aggregator.strategy
    m_before = [a.copy() for a in strat.m]
    real = srv.aggregator.aggregate

    def boom(*a, **k):
        strat.m[0] = strat.m[0] + 123.0  # a half-done update
        raise RuntimeError("aggregator down")
    srv.aggregator.aggregate = boom
    with pytest.raises(RuntimeError):
        _done(srv, "a", 3, [x + np.float32(0.03) for x in w0], {"num_samples": 4})
    np.testing.assert_array_equal(strat.m[0], m_before[0])
    srv.aggregator.aggregate = real
    up = [x + np.float32(0.03) for x in w0]
    c, _ = _done(srv, "a", 3, up, {"num_samples": 4})
    assert c["state"] == "RESP_ARY"
    g = ref.aggregate(g, [up], [4])
    np.testing.assert_allclose(srv.global_weights[0], g[0], rtol=1e-5, atol=1e-6)


def test_fedadam_server_resumes_with_its_moments(tmp_path):
    """A resumed FedAdam federation continues exactly like an uninterrupted one: the optimiser moments are part of the
    checkpoint (extension; the reference has neither a server optimiser nor a resume)."""
    w0 = S.init_weights(0)
    ups = {r: [x + np.float32(0.02 * r) for x in w0] for r in (1, 2, 3)}
    # uninterrupted
    a, _, _ = _server(tmp_path / "u", strategy="fedadam", max_rounds=5, aggregator="host")
    _ready(a, "c")
    for r in (1, 2, 3):
        _done(a, "c", r, ups[r], {"num_samples": 3})
    # interrupted after round 2
    b, _, cfg = _server(tmp_path / "i", strategy="fedadam", max_rounds=5, aggregator="host")
    _ready(b, "c")
    for r in (1, 2):
        _done(b, "c", r, ups[r], {"num_samples": 3})
    cfg.server.resume = True
    b2 = FLServer(cfg, S.init_weights(9), S, log=lambda s: None)
    assert b2.current_round == 3 and b2.model_version == 3 and b2.ready_clients == ["c"]
    assert b2.aggregator.strategy.m is not None
    c, _ = _done(b2, "c", 3, ups[3], {"num_samples": 3})
    assert c["state"] == "RESP_ARY"
    for x, y in zip(b2.global_weights, a.global_weights):
        np.testing.assert_allclose(x, y, rtol=1e-6, atol=1e-7)
