"""Mixed rounds: clients hosted on this box's GPUs AND off-box gRPC clients in ONE aggregation.

The reference knows one topology only: remote clients upload to the server, which averages on its CPU
(``fl_client.py:61-65,126-127`` -> ``fl_server.py:176-196``).  Here the gRPC service still terminates on the host of
rank 0 and still runs the reference's round state machine (``protocol/server.py``), but its aggregation step is this
bridge: the decoded 'D' payloads are staged host -> device into slots of rank 0's federation (K16) and take part in
the multi-GPU ``crackfl_fedagg`` launch as extra clients with their own sample counts; the server then answers its
clients (RESP_ARY / version poll) with the model that kernel produced.

Threading: ``aggregate`` runs on a gRPC worker thread (inside the server's lock) and blocks until the on-box round
loop (main thread of rank 0) has picked the updates up (``collect``), launched the fused kernel and handed the result
back (``deliver``).
"""
from __future__ import annotations

import threading
from typing import List, Optional, Sequence

import numpy as np

from ..io import weights as wio


class OnBoxBridge:
    kind = "onbox"

    def __init__(self, spec, remote_slots: int, timeout_s: float = 600.0):
        self.spec, self.max_clients, self.timeout_s = spec, int(remote_slots), float(timeout_s)
        self._cv = threading.Condition()
        self._pending = None
        self._result: Optional[List[np.ndarray]] = None
        self.rounds_bridged = 0

    # FedAdam state lives on the device side (runtime.OnBoxFederation); nothing to checkpoint here
    strategy = None

    # ---- gRPC worker thread (FLServer._aggregate_locked) -------------------------------------------------
    def aggregate(self, global_w, updates: Sequence[Sequence[np.ndarray]], num_samples=None) -> List[np.ndarray]:
        ns = list(num_samples) if num_samples is not None else [0.0] * len(updates)
        with self._cv:
            self._pending, self._result = (list(updates), ns), None
            self._cv.notify_all()
            if not self._cv.wait_for(lambda: self._result is not None, self.timeout_s):
                self._pending = None
                raise TimeoutError("the on-box round did not pick the remote updates up in time")
            out, self._result = self._result, None
            return out

    # ---- on-box round loop, rank 0 --------------------------------------------------------------------------
    def collect(self, fed, rnd: int) -> int:
        """Wait for the server to hand over this round's remote updates and stage them; returns how many."""
        with self._cv:
            if not self._cv.wait_for(lambda: self._pending is not None, self.timeout_s):
                return 0  # nobody reported: the round goes on with the on-box clients alone
            ups, ns = self._pending
        for u, n in zip(ups, ns):
            fed.stage_remote(wio.flatten(u, self.spec), n if n and n > 0 else 1.0)
        return len(ups)

    def deliver(self, fed):
        """After ``fed.aggregate``: give the waiting server thread the new global model."""
        with self._cv:
            if self._pending is None:
                return
            self._pending, self._result = None, fed.global_weights()
            self.rounds_bridged += 1
            self._cv.notify_all()
