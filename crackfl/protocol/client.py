"""Federated client driver: the round loop of ``fl_client.py:77-175``.

Flow (same message order as the reference): ``R`` ready -> ``P`` fetch global weights ->
``T`` training-started -> local fit -> loop { ``D`` train-done with weights ->
RESP_ACY: poll ``version P`` every ``poll_interval_s`` until NOT_WAIT/FIN |
RESP_ARY: ``T`` + fit next round with the weights in the reply | FIN: stop }.

Fixes over the reference: FIN is honoured on the poll path (quirk Q7: the reference
polls forever, ``fl_client.py:148``), ``fit`` returns the weights instead of handing
them over through ``./saved_weight/weights.pickle`` (Q8; the file is still written as a
checkpoint when ``client.write_weights_file``), gRPC errors surface as exceptions with
the phase in the message, payloads above the message cap are chunked.
"""
from __future__ import annotations

import os
import random
import time
from typing import Callable, Iterable, List, Optional

import numpy as np

from ..config import Config
from ..io import weights as wio
from . import messages as pb


def make_stub(cfg: Config, target: Optional[str] = None):
    """``run()`` (fl_client.py:178-183): insecure channel + stream-stream callable."""
    import grpc
    from .server import grpc_options
    channel = grpc.insecure_channel(target or cfg.transport.target, options=grpc_options(cfg))
    call = channel.stream_stream(f"/{cfg.transport.service_path}/{pb.METHOD}",
                                 request_serializer=pb.transportRequest.SerializeToString,
                                 response_deserializer=pb.transportResponse.FromString)
    return call, channel


# -- request builders (fl_client.py:14-74) ------------------------------------------------
def request_parameter():
    yield pb.transportRequest(update_req=pb.UpdateReq(type="P"))


def request_ready(name: str, current_round: int = 0):
    yield pb.transportRequest(ready_req=pb.ReadyReq(type="R", cname=name, state=pb.State.ON,
                                                    config=pb.to_config({"current_round": current_round})))


def request_training(name: str):
    yield pb.transportRequest(update_req=pb.UpdateReq(type="T", cname=name, state=pb.State.TRAINING))


def request_traindone(name: str, cr: int, payload: bytes, extra: dict, cap: int):
    n = max(1, (len(payload) + cap - 1) // cap)
    for i in range(n):
        cfgd = dict(extra)
        if n > 1:
            cfgd.update(chunk_index=i, chunk_count=n)
        yield pb.transportRequest(update_req=pb.UpdateReq(
            type="D", buffer_chunk=payload[i * cap:(i + 1) * cap], state=pb.State.TRAIN_DONE, cname=name,
            current_round=cr, config=pb.to_config(cfgd)))


def request_model_version(mv: int, cr: int, name: str = ""):
    cfgd = {"model_version": mv, "current_round": cr}
    if name:
        cfgd["cname"] = name  # extension: lets the server tick this client off when it answers FIN
    yield pb.transportRequest(version_req=pb.VersionReq(type="P", config=pb.to_config(cfgd)))


def get_file_chunks(filename: str, chunk_bytes: int):
    with open(filename, "rb") as f:
        i = 0
        while True:
            piece = f.read(chunk_bytes)
            if not piece and i > 0:
                return
            yield pb.transportRequest(update_req=pb.UpdateReq(
                type="L", buffer_chunk=piece, title=filename, file_len=len(piece),
                config=pb.to_config({"chunk_index": i})))
            i += 1
            if len(piece) < chunk_bytes:
                return


def _is_unavailable(e: Exception) -> bool:
    try:
        import grpc
        return isinstance(e, grpc.RpcError) and e.code() in (grpc.StatusCode.UNAVAILABLE, grpc.StatusCode.CANCELLED)
    except Exception:
        return False


def _collect_weights(responses: Iterable, field: str):
    """Concatenate (possibly chunked) ``buffer_chunk`` replies; returns (last_reply_submessage, bytes)."""
    last, parts = None, []
    for r in responses:
        last = getattr(r, field)
        parts.append(last.buffer_chunk)
    return last, b"".join(parts)


class FLClient:
    def __init__(self, cfg: Config, trainer, transport: Callable, sleep: Callable[[float], None] = time.sleep,
                 log: Callable[[str], None] = print):
        """``trainer`` implements ``fit(weights, round) -> (weights, metrics)`` (see
        ``crackfl.trainer.LocalTrainer``); ``transport`` is the stream-stream callable."""
        self.cfg, self.trainer, self.transport, self.sleep, self.log = cfg, trainer, transport, sleep, log
        self.name = cfg.client.name or f"client{random.randint(1, 100000)}"  # fl_client.py:26
        self.rounds_done = 0
        self.history: List[dict] = []
        self.final_weights: Optional[List[np.ndarray]] = None

    # ------------------------------------------------------------------
    def _loads(self, blob: bytes):
        return wio.loads_auto(blob, allow_pickle=self.cfg.transport.allow_pickle)

    def _fit(self, weights, cr: int):
        w, metrics = self.trainer.fit(weights, cr)
        self.history.append({"round": cr, **metrics})
        self.rounds_done += 1
        if self.cfg.client.write_weights_file and self.cfg.client.weights_path:
            wio.save_pickle_file(self.cfg.client.weights_path, w)  # client_fit_model.py:239-240
        return w, metrics

    def send_logs(self, path: str):
        for rep in self.transport(get_file_chunks(path, self.cfg.transport.chunk_mib << 20)):
            self.log(f"[{self.name}] delivered {rep.update_rep.title}")

    def upload_log_tree(self):
        root = self.cfg.client.log_root
        for d, _dirs, files in os.walk(root):
            for fn in files:
                self.send_logs(os.path.join(d, fn))

    # ------------------------------------------------------------------
    def run(self) -> dict:
        """``send_message`` (fl_client.py:77-175).  Returns a summary dict."""
        cfg = self.cfg
        info = None
        for rs in self.transport(request_ready(self.name)):
            c = pb.from_config(rs.ready_rep.config)
            if c.get("state") == "SW":
                info = {"cr": c["current_round"], "mtr": c["max_train_round"], "tmt": c.get("model_type", "unet"),
                        "mv": c["model_version"]}
                if c.get("cname"):
                    self.name = c["cname"]
        if info is None:
            self.log(f"[{self.name}] server replied CTW: admission window closed")
            return {"state": "CTW", "rounds": 0}

        _, blob = _collect_weights(self.transport(request_parameter()), "update_rep")
        weights = self._loads(blob)
        list(self.transport(request_training(self.name)))
        weights, metrics = self._fit(weights, info["cr"])

        cap = max(1, cfg.transport.max_message_mib - 1) << 20
        state = None
        while info["cr"] <= info["mtr"]:
            extra = {"num_samples": int(metrics.get("num_samples", 0))}
            for k in ("loss", "accuracy", "val_loss", "val_accuracy"):
                if k in metrics and metrics[k] is not None:
                    extra[k] = float(metrics[k])
            payload = wio.dumps(weights, cfg.transport.payload_format)
            rep, blob = _collect_weights(
                self.transport(request_traindone(self.name, info["cr"], payload, extra, cap)), "update_rep")
            if rep is None:
                raise RuntimeError(f"[{self.name}] no reply to train-done in round {info['cr']}")
            rc = pb.from_config(rep.config)
            state = rc.get("state")
            self.log(f"[{self.name}] round {info['cr']}: server state {state}")
            if state == "RESP_ACY":
                new = None
                while new is None:
                    self.sleep(cfg.client.poll_interval_s)  # fl_client.py:141
                    try:
                        vrep, vblob = _collect_weights(
                            self.transport(request_model_version(info["mv"], info["cr"], self.name)), "version_rep")
                    except Exception as e:
                        # a server that has shut down after the last round is the same news as FIN
                        if info["cr"] >= info["mtr"] and _is_unavailable(e):
                            self.log(f"[{self.name}] server gone after the last round: treated as FIN")
                            state, vrep = "FIN", None
                            break
                        raise
                    if vrep.state == pb.State.NOT_WAIT:
                        new = vblob
                    elif vrep.state == pb.State.FIN:
                        state = "FIN"
                        break
                vc = pb.from_config(vrep.config) if vrep is not None else {}
                info["cr"], info["mv"] = vc.get("current_round", info["cr"] + 1), vc.get("model_version", info["mv"])
                if state == "FIN":
                    break
                list(self.transport(request_training(self.name)))
                weights, metrics = self._fit(self._loads(new), info["cr"])
            elif state == "RESP_ARY":
                list(self.transport(request_training(self.name)))
                info["cr"], info["mv"] = rc["current_round"], rc["model_version"]
                weights, metrics = self._fit(self._loads(blob), info["cr"])
            elif state == "FIN":
                info["cr"], info["mv"] = rc["current_round"], rc["model_version"]
                break
            else:
                raise RuntimeError(f"[{self.name}] unexpected server state {state!r}")
        if cfg.client.upload_logs:
            self.upload_log_tree()
        # after FIN the aggregated final model can still be fetched with 'P'
        try:
            _, blob = _collect_weights(self.transport(request_parameter()), "update_rep")
            self.final_weights = self._loads(blob)
        except Exception:  # server may already be gone
            self.final_weights = None
        self.log(f"[{self.name}] all training finished after {self.rounds_done} local fits")
        return {"state": state or "FIN", "rounds": self.rounds_done, "name": self.name, "history": self.history}
