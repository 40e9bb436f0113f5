"""Single-process device aggregator: off-box (gRPC) updates staged H2D into arena slots and reduced
by the fused sm_100a kernel (SURVEY §8 K16 + K11/K12).  Drop-in for ``protocol.server.HostAggregator``.

The on-box multi-GPU path (one process per GPU, peer memory) is ``parallel/runtime.py``; both drive the
same kernel, ``crackfl_fedagg`` in ``csrc/fedagg.cu``.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch

from ..io import weights as wio
from ..ops import lib as L
from .fedavg import normalise_coefficients


class DeviceAggregator:
    kind = "device"

    def __init__(self, spec, strategy: str = "fedavg", payload: str = "fp32", max_clients: int = 64,
                 eta: float = 1e-2, beta1: float = 0.9, beta2: float = 0.99, tau: float = 1e-3,
                 device: Optional[int] = None):
        if payload not in ("fp32", "bf16_delta"):
            raise ValueError(payload)
        if max_clients > L.MAX_CLIENTS:
            raise ValueError(f"at most {L.MAX_CLIENTS} client slots per launch")
        self.lib = L.get_lib()
        self.spec, self.strategy_name, self.payload = spec, strategy, payload
        self.weighted = strategy in ("fedavg_weighted", "fedadam")
        self.server_opt = 1 if strategy == "fedadam" else 0
        self.eta, self.beta1, self.beta2, self.tau = eta, beta1, beta2, tau
        self.dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        n = spec.arena_numel
        self.n = n
        self.master = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.m = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.v = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.out = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.stage_f32 = torch.zeros((max_clients, n), dtype=torch.float32, device=self.dev)
        self.stage_bf16 = torch.zeros((max_clients, n), dtype=torch.bfloat16, device=self.dev) \
            if payload == "bf16_delta" else None
        mask = np.zeros(n, np.uint8)
        for t in spec.tensors:
            mask[t.offset:t.offset + t.numel] = 1 if t.trainable else 0
        self.trainable = torch.from_numpy(mask).to(self.dev)
        self.host = torch.empty((n,), dtype=torch.float32).pin_memory()
        self.max_clients = max_clients
        self.last_ms = 0.0
        self._master_loaded = False

    # FedAdam state mirrors the host strategy object so the server can checkpoint it
    @property
    def strategy(self):
        return self

    def state(self) -> dict:
        if not self.server_opt:
            return {}
        return {"m": wio.unflatten(self.m.cpu().numpy(), self.spec), "v": wio.unflatten(self.v.cpu().numpy(), self.spec)}

    def load_state(self, st: dict):
        if st.get("m") is not None:
            self.m.copy_(torch.from_numpy(wio.flatten(st["m"], self.spec)))
            self.v.copy_(torch.from_numpy(wio.flatten(st["v"], self.spec)))

    def aggregate(self, global_w: Sequence[np.ndarray], updates: Sequence[Sequence[np.ndarray]],
                  num_samples: Optional[Sequence[float]] = None) -> List[np.ndarray]:
        K = len(updates)
        if K < 1 or K > self.max_clients:
            raise ValueError(f"{K} updates, slots for {self.max_clients}")
        st = torch.cuda.current_stream(self.dev)
        with torch.cuda.device(self.dev):
            self.master.copy_(torch.from_numpy(wio.flatten(global_w, self.spec)))
            # K16: H2D staging.  Two pinned buffers in turn: flattening update k+1 on the host overlaps the copy of
            # update k, and only the copy that last read a buffer is waited for (the first version synchronised the
            # stream once per client).
            if not hasattr(self, "_pin"):
                self._pin = [self.host, torch.empty_like(self.host).pin_memory()]
                self._pin_ev = [None, None]
            for k, u in enumerate(updates):
                i = k & 1
                if self._pin_ev[i] is not None:
                    self._pin_ev[i].synchronize()
                np.copyto(self._pin[i].numpy(), wio.flatten(u, self.spec))
                self.stage_f32[k].copy_(self._pin[i], non_blocking=True)
                self._pin_ev[i] = torch.cuda.Event()
                self._pin_ev[i].record(st)
                if self.payload == "bf16_delta":
                    L.check(self.lib.crackfl_publish_delta(self.stage_f32[k].data_ptr(), self.master.data_ptr(),
                                                           self.stage_bf16[k].data_ptr(), self.n, st.cuda_stream))
            coef = normalise_coefficients(K, num_samples if (self.weighted and num_samples is not None) else None)
            p = L.FedAggParams()
            src = self.stage_bf16 if self.payload == "bf16_delta" else self.stage_f32
            for k in range(K):
                p.client[k] = src[k].data_ptr()
                p.coef[k] = float(coef[k])
            p.out[0] = self.out.data_ptr()
            p.num_clients, p.num_out = K, 1
            p.payload_bf16 = 1 if self.payload == "bf16_delta" else 0
            p.out_bf16 = 0
            p.delta = 1 if self.payload == "bf16_delta" else 0
            p.server_opt = self.server_opt
            p.eta, p.beta1, p.beta2, p.tau = self.eta, self.beta1, self.beta2, self.tau
            p.master, p.m, p.v = self.master.data_ptr(), self.m.data_ptr(), self.v.data_ptr()
            p.trainable = self.trainable.data_ptr()
            p.mc_out = 0
            p.begin, p.end = 0, self.n
            p.rank, p.world, p.epoch = 0, 1, 1
            p.error_flag, p.local_sync = 0, 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            L.check(self.lib.crackfl_fedagg(C.byref(p), 148 * 4, st.cuda_stream), "fedagg")
            e1.record(st)
            st.synchronize()
            self.last_ms = e0.elapsed_time(e1)
            return wio.unflatten(self.out.cpu().numpy(), self.spec)
