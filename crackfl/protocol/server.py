"""FedAvg coordinator: round state machine + gRPC service.

Capability parity with ``fl_server.py`` (all of it), re-designed as a single-owner state
machine (one lock, injectable clock) instead of module globals mutated from ten gRPC
worker threads with no lock (``fl_server.py:14-20``, quirk Q11).

Message vocabulary and state strings are the reference's:

==========  =========================================  ==================================
request     reference handler                           reply
==========  =========================================  ==================================
ready  'R'  ``ready_client``   fl_server.py:45-81       config{state SW|CTW, model_version,
                                                        current_round, max_train_round, model_type}
update 'P'  ``send_parameter`` fl_server.py:23-24       update_rep{type P, buffer_chunk, title "parameters"}
update 'T'  ack               fl_server.py:162-169      update_rep{type T}
update 'L'  ``save_chunks_to_file`` :84-89,170-175      update_rep{type L, title}
update 'D'  ``manage_rounds`` fl_server.py:107-135      update_rep{type D, config{state RESP_ACY|RESP_ARY|FIN,...}}
version 'P' ``version_check`` fl_server.py:138-149      version_rep{state WAIT|NOT_WAIT|FIN, ...}
==========  =========================================  ==================================

Deliberate fixes: Q1 per-round mean (updates cleared each round), Q2 the
broadcast carries the *aggregated* weights, Q7 FIN is a real enum label, Q11 locking,
Q12 correctly spelt send limit, Q13 directories created, Q14 model_type "unet".
Extensions: sample-count weighting, FedAdam, quorum + round deadline, server-issued
unique client ids, chunked payloads above the message cap, resume from checkpoint.
"""
from __future__ import annotations

import json
import os
import threading
import time
from concurrent import futures
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

from ..config import Config
from ..io import checkpoint as ckpt
from ..io import weights as wio
from ..parallel.fedavg import make_strategy
from . import messages as pb

SW, CTW = "SW", "CTW"
RESP_ACY, RESP_ARY, FIN = "RESP_ACY", "RESP_ARY", "FIN"
MODEL_TYPE_ALIASES = {"mobilenet_v2": "unet", "unet": "unet"}


class HostAggregator:
    """NumPy aggregation (the reference's place for it: server CPU, ``fl_server.py:92-102``)."""

    kind = "host"

    def __init__(self, strategy):
        self.strategy = strategy

    def aggregate(self, global_w, updates, num_samples=None):
        return self.strategy.aggregate(global_w, updates, num_samples)


@dataclass
class FaultInjector:
    """Test hook: drop / delay / duplicate inbound requests by type tag."""

    drop: Dict[str, int] = field(default_factory=dict)  # tag -> how many to drop
    delay_s: Dict[str, float] = field(default_factory=dict)
    duplicate: Dict[str, int] = field(default_factory=dict)

    def on_request(self, tag: str, sleep: Callable[[float], None]) -> str:
        if self.delay_s.get(tag):
            sleep(self.delay_s[tag])
        if self.drop.get(tag, 0) > 0:
            self.drop[tag] -= 1
            return "drop"
        if self.duplicate.get(tag, 0) > 0:
            self.duplicate[tag] -= 1
            return "duplicate"
        return "ok"


class FLServer:
    def __init__(self, cfg: Config, initial_weights: Sequence[np.ndarray], spec=None, aggregator=None,
                 clock: Callable[[], float] = time.monotonic, sleep: Callable[[float], None] = time.sleep,
                 fault: Optional[FaultInjector] = None, log: Callable[[str], None] = print,
                 eval_fn: Optional[Callable[[List[np.ndarray]], dict]] = None):
        from ..models.spec import REFERENCE_SPEC
        self.cfg = cfg
        self.spec = spec or REFERENCE_SPEC
        self.clock, self.sleep, self.fault, self.log = clock, sleep, fault, log
        # optional server-side evaluation of every new global model: the reference's dead ``trainNextRound`` hook
        # (fl_server.py:27-37, call site commented out at :65); returns e.g. {'loss':..., 'accuracy':...}
        self.eval_fn = eval_fn
        self.lock = threading.RLock()
        s = cfg.server
        if aggregator is None:
            kw = {}
            if s.strategy == "fedadam":
                kw = dict(eta=s.fedadam_eta, beta1=s.fedadam_beta1, beta2=s.fedadam_beta2, tau=s.fedadam_tau,
                          trainable_mask=[t.trainable for t in self.spec.tensors])
            aggregator = HostAggregator(make_strategy(s.strategy, **kw))
        self.aggregator = aggregator
        # ---- state (reference globals, fl_server.py:14-20) ----
        self.global_weights: List[np.ndarray] = [np.asarray(w, np.float32) for w in initial_weights]
        self.ready_clients: List[str] = []  # readyClientSids
        self.round_reporters: List[str] = []  # currentRoundClientUpdates
        self.round_updates: List[List[np.ndarray]] = []  # received_parameters (cleared every round: fix Q1)
        self.round_samples: List[float] = []
        self.round_metrics: List[dict] = []
        self.current_round = 0  # ALL_CURRENT_ROUND
        self.model_version = s.initial_model_version  # MODEL_VERSION
        self.window_deadline: Optional[float] = None  # CHECK_TRAIN_TIMER / CHECK_CLIENT_TRAINING
        self.round_started_at: Optional[float] = None
        self.finished = False
        self.history: List[dict] = []
        self._payload_cache = {}
        self._chunks: Dict[str, tuple] = {}  # cname -> (chunk_count, {index: bytes})
        self._next_id = 0
        self.final_clients: List[str] = []  # reporters of the last round: each must be told FIN before the server stops
        self.fin_answered: set = set()
        self.finished_at: Optional[float] = None
        if s.resume and ckpt.has_checkpoint(s.weights_dir):
            self._resume()

    # ------------------------------------------------------------------ helpers
    def _window_closed(self) -> bool:
        return self.window_deadline is not None and self.clock() >= self.window_deadline

    def serialized_weights(self) -> bytes:
        """``send_parameter`` (fl_server.py:23-24); serialised once per model version."""
        fmt = self.cfg.transport.payload_format
        key = (self.model_version, fmt)
        if key not in self._payload_cache:
            self._payload_cache = {key: wio.dumps(self.global_weights, fmt)}
        return self._payload_cache[key]

    def _target_count(self) -> int:
        return len(self.ready_clients)

    # ------------------------------------------------------------------ handlers
    def ready_client(self, name: str, config: dict) -> dict:
        s = self.cfg.server
        with self.lock:
            if self.window_deadline is None:
                self.window_deadline = self.clock() + s.admission_window_s
            if self._window_closed() or self.finished:
                return {"state": CTW}
        if s.ready_delay_s > 0:
            self.sleep(s.ready_delay_s)  # fl_server.py:56 (outside the lock)
        with self.lock:
            assigned = name
            if not name or (name in self.ready_clients and config.get("rejoin", 0) != 1):
                # reference names are client{randint(1,100000)} and can collide (fl_client.py:26)
                self._next_id += 1
                assigned = f"{name or 'client'}#{self._next_id}"
            cap = getattr(self.aggregator, "max_clients", 0) or 0
            if cap and assigned not in self.ready_clients and len(self.ready_clients) >= cap:
                return {"state": CTW}  # the device aggregator has a fixed number of staging slots
            if assigned not in self.ready_clients:
                self.ready_clients.append(assigned)
            out = {}
            if int(config.get("current_round", 0) or 0) == 0:
                # the reference does ``ALL_CURRENT_ROUND = 1`` on every such request (fl_server.py:61-63); with rounds
                # that last seconds a client can be admitted AFTER the first aggregation, and resetting the counter then
                # would strand everybody (stale-drop + WAIT forever).  Only the very first ready starts round 1; a later
                # joiner is handed the live round and model version.
                if self.current_round == 0:
                    self.current_round = 1
                if self.round_started_at is None:
                    self.round_started_at = self.clock()
                out = {"state": SW, "model_version": self.model_version, "current_round": self.current_round,
                       "max_train_round": s.max_rounds, "model_type": s.model_type, "cname": assigned,
                       "payload_format": self.cfg.transport.payload_format}
            return out

    def _aggregate_locked(self) -> str:
        s = self.cfg.server
        t0 = time.perf_counter()
        ns = self.round_samples if all(n > 0 for n in self.round_samples) else None
        self.global_weights = [np.asarray(w, np.float32) for w in
                               self.aggregator.aggregate(self.global_weights, self.round_updates, ns)]
        agg_ms = (time.perf_counter() - t0) * 1e3
        rec = {"round": self.current_round, "clients": list(self.round_reporters), "num_clients": len(self.round_updates),
               "samples": list(self.round_samples), "aggregate_ms": agg_ms, "model_version": self.model_version + 1,
               "aggregator": getattr(self.aggregator, "kind", "?"), "client_metrics": list(self.round_metrics)}
        if self.eval_fn is not None:
            try:
                rec["global_eval"] = self.eval_fn(self.global_weights)
                self.log(f"[server] round {self.current_round} global model: {rec['global_eval']}")
            except Exception as e:  # evaluation must never stall a round
                rec["global_eval_error"] = repr(e)
        self.history.append(rec)
        finished_round = self.current_round
        # clients that did not report are no longer waited for in later rounds
        self.ready_clients = [c for c in self.ready_clients if c in self.round_reporters] or self.ready_clients
        self.round_reporters, self.round_updates, self.round_samples, self.round_metrics = [], [], [], []
        self.current_round += 1
        self.model_version += 1
        self.round_started_at = self.clock()
        self._persist(rec)
        if finished_round >= s.max_rounds:
            self.finished = True
            self.final_clients = list(rec["clients"])
            self.finished_at = self.clock()
            self.log(f"[server] all {s.max_rounds} rounds done -> FIN")
            return FIN
        self.log(f"[server] round {finished_round} aggregated ({rec['num_clients']} clients, {agg_ms:.2f} ms) -> RESP_ARY")
        return RESP_ARY

    def _maybe_timeout_locked(self):
        s = self.cfg.server
        if s.round_timeout_s <= 0 or self.finished or not self.round_updates or self.round_started_at is None:
            return
        if self.clock() - self.round_started_at >= s.round_timeout_s and \
                len(self.round_updates) >= max(1, s.min_fit_clients):
            self.log(f"[server] round {self.current_round} deadline: aggregating {len(self.round_updates)} of "
                     f"{self._target_count()} clients")
            self._aggregate_guarded(self.round_reporters[-1])

    def manage_rounds(self, cname: str, client_round: int, weights: List[np.ndarray], num_samples: float = 0.0,
                      metrics: Optional[dict] = None) -> str:
        with self.lock:
            self._maybe_timeout_locked()
            if self.finished or client_round != self.current_round:
                return FIN if self.finished else RESP_ACY  # stale update dropped (fl_server.py:134-135)
            if cname in self.round_reporters:
                return RESP_ACY  # duplicate report
            weights = self._validated(cname, weights)
            self.round_reporters.append(cname)
            self.round_updates.append(weights)
            self.round_samples.append(float(num_samples))
            if metrics:
                self.round_metrics.append({"cname": cname, **metrics})
            target = self._target_count()
            s = self.cfg.server
            if s.round_timeout_s <= 0 and s.min_fit_clients > 0:
                target = min(target, s.min_fit_clients)
            if len(self.round_updates) >= target:
                return self._aggregate_guarded(cname)
            return RESP_ACY

    def _validated(self, cname: str, weights) -> List[np.ndarray]:
        """Reject an update BEFORE it is queued: tensor count, exact shapes (NumPy broadcasting would silently accept
        ``(1,)`` or ``()``), float dtype, finite values.  Raises ValueError; nothing is stored."""
        if len(weights) != self.spec.num_tensors:
            raise ValueError(f"update from {cname} has {len(weights)} tensors, expected {self.spec.num_tensors}")
        out = []
        for t, w in zip(self.spec.tensors, weights):
            a = np.asarray(w)
            if tuple(a.shape) != tuple(t.shape):
                raise ValueError(f"update from {cname}: tensor {t.name} has shape {tuple(a.shape)}, expected {tuple(t.shape)}")
            if a.dtype.kind != "f":
                raise ValueError(f"update from {cname}: tensor {t.name} has dtype {a.dtype}, expected a float type")
            if self.cfg.server.check_finite and not np.isfinite(a).all():
                raise ValueError(f"update from {cname}: tensor {t.name} holds non-finite values")
            out.append(a)
        return out

    def _aggregate_guarded(self, cname: str) -> str:
        """Aggregate; if the aggregator raises, evict the update that triggered it (and restore the optimiser state)
        so that one bad client cannot wedge every later 'D' of the round on the same exception."""
        strat = getattr(self.aggregator, "strategy", None)
        snap = strat.state() if strat is not None and hasattr(strat, "state") else None
        if snap and snap.get("m") is not None:
            # deep copy: the moments are LISTS of per-tensor arrays (different shapes) that aggregate() re-binds in place
            snap = {k: ([np.array(a, copy=True) for a in v] if isinstance(v, (list, tuple)) else np.array(v, copy=True))
                    for k, v in snap.items() if v is not None}
        else:
            snap = None
        try:
            return self._aggregate_locked()
        except Exception:
            if cname in self.round_reporters:
                i = self.round_reporters.index(cname)
                for lst in (self.round_reporters, self.round_updates, self.round_samples):
                    del lst[i]
                self.round_metrics = [m for m in self.round_metrics if m.get("cname") != cname]
            if cname in self.ready_clients and len(self.ready_clients) > 1:
                self.ready_clients.remove(cname)  # not waited for any more
            if snap is not None:
                strat.load_state(snap)
            self.log(f"[server] aggregation failed on the update from {cname}: evicted")
            raise

    def version_check(self, client_version: int, client_round: int):
        with self.lock:
            self._maybe_timeout_locked()
            if self.model_version == client_version:
                return pb.State.WAIT, {}
            cfgd = {"current_round": self.current_round, "model_version": self.model_version}
            if client_round >= self.cfg.server.max_rounds:
                return pb.State.FIN, cfgd
            return pb.State.NOT_WAIT, cfgd

    def save_log_chunk(self, chunk: bytes, title: str, append: bool) -> str:
        """'L' sink.  Reference strips the 11-char './send_logs' prefix and writes under '.'
        (fl_server.py:84-89); here the path is confined to ``cfg.server.log_dir``."""
        rel = title[11:] if title.startswith("./send_logs") else title
        rel = os.path.normpath("/" + rel.replace("\\", "/")).lstrip("/")
        root = os.path.abspath(self.cfg.server.log_dir)
        path = os.path.abspath(os.path.join(root, rel))
        if not path.startswith(root + os.sep):
            raise ValueError("log path escapes the log directory")
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "ab" if append else "wb") as f:
            f.write(chunk)
        return path

    # ------------------------------------------------------------------ dispatch
    def handle(self, req) -> list:
        """One request -> list of responses (exactly one, except chunked weight replies)."""
        R = pb.transportResponse
        if req.ready_req.type == "R":
            if self._fault("R") == "drop":
                return []
            cfgd = self.ready_client(req.ready_req.cname, pb.from_config(req.ready_req.config))
            return [R(ready_rep=pb.ReadyRep(config=pb.to_config(cfgd)))]
        u = req.update_req
        if u.type == "P":
            return self._weight_replies(lambda chunk, cfgd: R(update_rep=pb.UpdateRep(
                type="P", buffer_chunk=chunk, title="parameters", config=pb.to_config(cfgd))), {})
        if u.type == "T":
            return [R(update_rep=pb.UpdateRep(type="T"))]
        if u.type == "L":
            ucfg = pb.from_config(u.config)
            self.save_log_chunk(u.buffer_chunk, u.title, append=int(ucfg.get("chunk_index", 0) or 0) > 0)
            return [R(update_rep=pb.UpdateRep(type="L", title=u.title))]
        if u.type == "D":
            f = self._fault("D")
            if f == "drop":
                return []
            ucfg = pb.from_config(u.config)
            payload = self._assemble_chunks(u.cname, u.buffer_chunk, ucfg)
            if payload is None:
                return [R(update_rep=pb.UpdateRep(type="D", config=pb.to_config({"state": "CHUNK"})))]
            weights = wio.loads_auto(payload, allow_pickle=self.cfg.transport.allow_pickle)
            metrics = {k: ucfg[k] for k in ("loss", "accuracy", "val_loss", "val_accuracy") if k in ucfg}
            state = self.manage_rounds(u.cname, u.current_round, weights, float(ucfg.get("num_samples", 0) or 0), metrics)
            if f == "duplicate":
                self.manage_rounds(u.cname, u.current_round, weights)
            cfgd = {"model_version": self.model_version, "current_round": self.current_round, "state": state}
            if state == FIN:
                self._note_fin(u.cname)
            if state == RESP_ARY:
                return self._weight_replies(lambda chunk, c: R(update_rep=pb.UpdateRep(
                    type="D", buffer_chunk=chunk, config=pb.to_config(c))), cfgd)
            return [R(update_rep=pb.UpdateRep(type="D", config=pb.to_config(cfgd)))]
        if req.version_req.type == "P":
            vc = pb.from_config(req.version_req.config)
            st, cfgd = self.version_check(int(vc.get("model_version", 0) or 0), int(vc.get("current_round", 0) or 0))
            if st == pb.State.FIN:
                self._note_fin(str(vc.get("cname", "") or ""))
            if st == pb.State.NOT_WAIT:
                return self._weight_replies(lambda chunk, c: R(version_rep=pb.VersionRep(
                    state=st, buffer_chunk=chunk, config=pb.to_config(c))), cfgd)
            return [R(version_rep=pb.VersionRep(state=st, config=pb.to_config(cfgd)))]
        raise ValueError("unrecognised request")

    def _note_fin(self, cname: str):
        with self.lock:
            self.fin_answered.add(cname or f"?{len(self.fin_answered)}")

    def all_fin_delivered(self) -> bool:
        """True once every reporter of the final round has been sent FIN (D reply or version poll).  Polls from clients
        that do not name themselves (the reference's version request has no cname) are counted, not matched."""
        with self.lock:
            if not self.finished:
                return False
            named = {c for c in self.fin_answered if not c.startswith("?")}
            anon = len(self.fin_answered) - len(named)
            missing = [c for c in self.final_clients if c not in named]
            return len(missing) <= anon

    def _fault(self, tag: str) -> str:
        return self.fault.on_request(tag, self.sleep) if self.fault else "ok"

    def _weight_replies(self, make, cfgd: dict) -> list:
        blob = self.serialized_weights()
        cap = max(1, self.cfg.transport.max_message_mib - 1) * (1 << 20)
        if len(blob) <= cap:
            return [make(blob, cfgd)]
        n = (len(blob) + cap - 1) // cap
        return [make(blob[i * cap:(i + 1) * cap], {**cfgd, "chunk_index": i, "chunk_count": n}) for i in range(n)]

    MAX_CHUNKS = 4096

    def _assemble_chunks(self, cname: str, chunk: bytes, ucfg: dict) -> Optional[bytes]:
        """Chunked 'D' payloads (extension; the reference sends one message, fl_client.py:61-65).  Returns the whole payload
        once every index 0..n-1 has arrived (any order, a repeated index overwrites), None before that.  A malformed
        header is a ValueError for that client only; a different chunk_count, or an unchunked payload, from the same
        client supersedes what it had half-sent."""
        n = int(ucfg.get("chunk_count", 0) or 0)
        if n <= 1:
            with self.lock:
                self._chunks.pop(cname, None)
            return chunk
        i = int(ucfg.get("chunk_index", 0) or 0)
        if n > self.MAX_CHUNKS or not 0 <= i < n:
            raise ValueError(f"update from {cname}: chunk {i} of {n} is not acceptable")
        with self.lock:
            ent = self._chunks.get(cname)
            if ent is None or ent[0] != n:
                ent = (n, {})
                self._chunks[cname] = ent
            ent[1][i] = chunk
            if len(ent[1]) < n:
                return None
            del self._chunks[cname]
        return b"".join(ent[1][k] for k in range(n))

    # gRPC servicer entry point (``TransportService.transport`` fl_server.py:209-212)
    def transport(self, request_iterator, context=None):
        for req in request_iterator:
            for rep in self.handle(req):
                yield rep

    # ------------------------------------------------------------------ persistence
    def _persist(self, rec: dict):
        d = self.cfg.server.weights_dir
        if not d:
            return
        os.makedirs(d, exist_ok=True)
        wio.save_pickle_file(os.path.join(d, "weights.pickle"), self.global_weights)  # fl_server.py:104-105
        strat = getattr(self.aggregator, "strategy", None)
        extra = {}
        st = strat.state() if strat is not None else {}
        if st.get("m") is not None:
            extra = {"fedadam_m": st["m"], "fedadam_v": st["v"]}
        ckpt.save_checkpoint(d, self.spec, self.global_weights, extra_arenas=extra, meta={
            "current_round": self.current_round, "model_version": self.model_version, "finished": self.finished,
            "ready_clients": self.ready_clients, "strategy": self.cfg.server.strategy})
        with open(os.path.join(d, "rounds.jsonl"), "a") as f:
            f.write(json.dumps(rec) + "\n")

    def _resume(self):
        st = ckpt.load_checkpoint(self.cfg.server.weights_dir, self.spec)
        self.global_weights = st["weights"]
        m = st["meta"]
        self.current_round = int(m.get("current_round", 0))
        self.model_version = int(m.get("model_version", self.model_version))
        self.finished = bool(m.get("finished", False))
        self.ready_clients = list(m.get("ready_clients", []))
        strat = getattr(self.aggregator, "strategy", None)
        if strat is not None and "fedadam_m" in st["arenas"]:
            strat.load_state({"m": st["arenas"]["fedadam_m"], "v": st["arenas"]["fedadam_v"]})
        if self.ready_clients:
            self.window_deadline = self.clock()  # window already closed for a resumed federation
        self.log(f"[server] resumed at round {self.current_round}, model_version {self.model_version}")


# ---------------------------------------------------------------------- gRPC plumbing
def grpc_options(cfg: Config):
    n = cfg.transport.max_message_mib * 1024 * 1024
    return [("grpc.max_receive_message_length", n), ("grpc.max_send_message_length", n)]


def make_grpc_server(server: FLServer, bind: Optional[str] = None):
    """``serve()`` (fl_server.py:214-226).  Returns ``(grpc_server, bound_port)``; caller starts/stops it."""
    import grpc
    cfg = server.cfg
    gs = grpc.server(futures.ThreadPoolExecutor(max_workers=cfg.transport.workers), options=grpc_options(cfg))
    handler = grpc.stream_stream_rpc_method_handler(
        server.transport, request_deserializer=pb.transportRequest.FromString,
        response_serializer=pb.transportResponse.SerializeToString)
    gs.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(cfg.transport.service_path, {pb.METHOD: handler}),))
    port = gs.add_insecure_port(bind or cfg.transport.bind)
    return gs, port


def serve_forever(server: FLServer):
    gs, port = make_grpc_server(server)
    gs.start()
    server.log(f"[server] listening on {server.cfg.transport.bind} (port {port})")
    # The reference server never stops (fl_server.py:222-226).  This one does, but only after every client of the last
    # round has been told FIN: all but the last reporter learn it from a ``version P`` poll, which they send only every
    # client.poll_interval_s (20 s by default), so stopping right after the aggregation would hand them a dead channel.
    grace = server.cfg.server.fin_grace_s or (server.cfg.client.poll_interval_s + 5.0)
    try:
        while not server.finished:
            time.sleep(0.2)
        t_end = time.monotonic() + grace
        while not server.all_fin_delivered() and time.monotonic() < t_end:
            time.sleep(0.2)
        time.sleep(1.0)  # let the last replies (and a final 'P' fetch of the aggregated model) drain
    except KeyboardInterrupt:
        pass
    gs.stop(2.0)
