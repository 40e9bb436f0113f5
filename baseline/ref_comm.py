#!/usr/bin/env python
"""R-comm baseline: the reference's OWN gRPC + pickle + NumPy round path, timed.

Runs the UNMODIFIED ``/root/reference/fl_server.py`` and ``fl_client.py`` as subprocesses (via ``runpy``) against
shim modules for what the reference imports but does not ship or cannot import on this image:

* ``transport_pb2`` / ``transport_pb2_grpc``  - reconstructed descriptors (crackfl/protocol/messages.py);
* ``model_evaluate``                          - stub global model: ``get_weights`` returns ``pickle.dumps(list of
  float32 arrays)`` of the requested size, re-reading ``./server_weights/weights.pickle`` so that FedAvg feeds back;
* ``client_fit_model``                        - stub ``learning_fit`` that SKIPS the Keras fit (TensorFlow is not
  installable) and just hands the received weights (+ a constant) back through ``./saved_weight/weights.pickle``.

So what is measured is exactly the reference's communication + aggregation + round-barrier logic (its code, its
message flow, its pickling, its NumPy mean, its 5 s / 10 s / 20 s sleeps) with zero training time.  Label:
"reference gRPC + pickle + NumPy path (reference code; proto and missing modules reconstructed)".

Two known reference bugs are visible in the timings and are worked around only by the harness' timeout:
the final-round straggler polls forever (fl_client.py:148, quirk Q7) and ``received_parameters`` is never cleared
(fl_server.py:127-130, quirk Q1: round r sums r*N arrays).
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("CRACKFL_REFERENCE_DIR", "/root/reference")

SHIMS = {
    "transport_pb2.py": '''
import os, time as _time
from crackfl.protocol.messages import *  # noqa: F401,F403  (Scalar, transportRequest, ..., State)
_mode = os.environ.get("REF_SLEEP_MODE", "shipped")
if _mode == "nopoll":
    # keep the registration sleeps (5 s per 'R', 10 s window: they are what makes every client register before the
    # first one reports, fl_server.py:42,56) but turn the client's 20 s version poll (fl_client.py:141) into a busy poll
    _orig = _time.sleep
    def _sleep(s):
        _orig(0.002 if s == 20 else s)
    _time.sleep = _sleep
''',
    "transport_pb2_grpc.py": '''
from crackfl.compat.transport_pb2_grpc import *  # noqa: F401,F403
''',
    "model_evaluate.py": '''
import os, pickle, numpy as np
def _initial():
    mb = float(os.environ.get("REF_PAYLOAD_MB", "8.233"))
    n = int(mb * 1e6 / 4)
    k = 112
    rng = np.random.default_rng(0)
    return [rng.standard_normal(n // k).astype(np.float32) for _ in range(k)]
class evaluate_LocalModel(object):
    def __init__(self, batch_size, img_size, class_names):
        pass
    def buildGlobalModel(self, channels, lr):
        os.makedirs("./server_weights", exist_ok=True)
        return {"weights": _initial()}
    def get_weights(self, model):
        p = "./server_weights/weights.pickle"
        if os.path.exists(p):
            with open(p, "rb") as f:
                return f.read()
        return pickle.dumps(model["weights"])
    def train_model_tosave(self, model):
        return {"loss": 0.0, "accuracy": 0.0}
''',
    "client_fit_model.py": '''
import os, pickle, time
class learning_fit(object):
    """Stub of client_fit_model.learning_fit: no Keras fit (TensorFlow unavailable); same hand-off file."""
    def __init__(self, mt, ec, bs, pr, cr):
        self.params = pickle.loads(pr)
    def manage_train(self, params=None, cr=None):
        print(f"### Model Training - Round: {cr} ###", flush=True)
        if params is not None:
            params = pickle.loads(params)
        else:
            params = self.params
        params = [p + 0.001 for p in params]
        os.makedirs("./saved_weight", exist_ok=True)
        with open("./saved_weight/weights.pickle", "wb") as fw:
            pickle.dump(params, fw)
''',
}

RUNNER = "import runpy,sys; sys.argv=[sys.argv[1]]; runpy.run_path(sys.argv[0], run_name='__main__')"


def run_once(num_clients: int, payload_mb: float, sleep_mode: str, timeout_s: float) -> dict:
    work = tempfile.mkdtemp(prefix="refcomm_")
    shim = os.path.join(work, "shim")
    os.makedirs(shim)
    for fn, src in SHIMS.items():
        with open(os.path.join(shim, fn), "w") as f:
            f.write(src)
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([shim, REPO]), REF_PAYLOAD_MB=str(payload_mb),
               REF_SLEEP_MODE=sleep_mode, PYTHONUNBUFFERED="1")
    sdir = os.path.join(work, "server")
    os.makedirs(os.path.join(sdir, "server_weights"))
    slog = open(os.path.join(work, "server.log"), "w")
    server = subprocess.Popen([sys.executable, "-c", RUNNER, os.path.join(REF, "fl_server.py")], cwd=sdir, env=env,
                              stdout=slog, stderr=subprocess.STDOUT)
    time.sleep(1.5)
    t0 = time.perf_counter()
    clients, logs = [], []
    for i in range(num_clients):
        cdir = os.path.join(work, f"client{i}")
        os.makedirs(os.path.join(cdir, "saved_weight"))
        lg = open(os.path.join(work, f"client{i}.log"), "w")
        logs.append(lg)
        clients.append(subprocess.Popen([sys.executable, "-c", RUNNER, os.path.join(REF, "fl_client.py")], cwd=cdir, env=env,
                                        stdout=lg, stderr=subprocess.STDOUT))
    # done when the server has logged FIN (the last reporter of round 5 got it); stragglers hang by design (Q7)
    fin_t, first_round_t = None, None
    while time.perf_counter() - t0 < timeout_s:
        txt = open(os.path.join(work, "server.log")).read()
        if first_round_t is None and "RESP_ARY" in txt:
            first_round_t = time.perf_counter() - t0
        if "Current state: FIN" in txt:
            fin_t = time.perf_counter() - t0
            break
        if server.poll() is not None:
            break
        time.sleep(0.02)
    finished = sum(1 for c in clients if c.poll() is not None)
    for p in clients + [server]:
        if p.poll() is None:
            p.kill()
    for lg in logs + [slog]:
        lg.close()
    stxt = open(os.path.join(work, "server.log")).read()
    per_round = (fin_t - first_round_t) / 4.0 if (fin_t and first_round_t) else None
    rec = {"clients": num_clients, "payload_mb": payload_mb, "sleep_mode": sleep_mode, "rounds": 5,
           "t_first_round_s": first_round_t, "t_all_rounds_s": fin_t, "per_round_s_rounds_2_to_5": per_round,
           "server_errors": stxt.count("Traceback"),
           "clients_exited_cleanly": finished, "label": "reference gRPC + pickle + NumPy path (reference code; proto and "
           "missing modules reconstructed; local fit stubbed out)"}
    if fin_t is None:
        rec["server_log_tail"] = stxt[-600:]
    shutil.rmtree(work, ignore_errors=True)
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clients", default="2")
    ap.add_argument("--payload-mb", default="8.233")
    ap.add_argument("--sleep-modes", default="nopoll,shipped",
                    help="shipped = 5/10/20 s sleeps as in the reference; nopoll = 20 s version poll turned into a busy poll")
    ap.add_argument("--timeout", type=float, default=400)
    ap.add_argument("--out", default="baseline/ref_comm_results.json")
    a = ap.parse_args()
    out = []
    for sc in a.sleep_modes.split(","):
        for n in [int(x) for x in a.clients.split(",")]:
            for mb in [float(x) for x in a.payload_mb.split(",")]:
                rec = run_once(n, mb, sc, a.timeout)
                print(json.dumps(rec), flush=True)
                out.append(rec)
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
