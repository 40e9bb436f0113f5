"""Server-side aggregation rules (host oracles + strategy objects).

``fedavg_mean`` restates the arithmetic of the reference's ``updateWeight``
(``fl_server.py:92-102``): element-wise sum of every client's list of arrays divided by
the number of reporting clients.  The reference never clears ``received_parameters``
(quirk Q1, ``fl_server.py:127-130``) so from round 2 on it averages stale updates too;
that is fixed here: a round only averages that round's updates.

``fedavg_weighted`` (sample-count weighting) and ``FedAdam`` are north-star extensions
(BASELINE.json configs 3 and 4); the reference has neither.

These NumPy functions are the numerical oracle for the fused sm_100a aggregation
kernel in ``csrc/fedagg.cu``; the runtime on a GPU box never dispatches to them.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

WeightList = List[np.ndarray]


def fedavg_mean(updates: Sequence[WeightList]) -> WeightList:
    """Unweighted mean over clients, per tensor (reference behaviour for a single round)."""
    if not updates:
        raise ValueError("no client updates to aggregate")
    n = len(updates)
    acc = [np.array(w, dtype=np.float32, copy=True) for w in updates[0]]
    for wcl in updates[1:]:
        if len(wcl) != len(acc):
            raise ValueError("clients disagree on the number of tensors")
        for i, wc in enumerate(wcl):
            acc[i] = acc[i] + np.asarray(wc, dtype=np.float32)
    return [a / np.float32(n) for a in acc]


def fedavg_weighted(updates: Sequence[WeightList], num_samples: Sequence[float]) -> WeightList:
    """Sample-count weighted mean: sum_k n_k w_k / sum_k n_k (McMahan et al. FedAvg)."""
    if len(updates) != len(num_samples) or not updates:
        raise ValueError("need one sample count per update")
    tot = float(sum(num_samples))
    if tot <= 0:
        raise ValueError("total sample count must be positive")
    coef = [np.float32(n / tot) for n in num_samples]
    acc = [np.asarray(w, np.float32) * coef[0] for w in updates[0]]
    for c, wcl in zip(coef[1:], updates[1:]):
        for i, wc in enumerate(wcl):
            acc[i] = acc[i] + np.asarray(wc, np.float32) * c
    return acc


def normalise_coefficients(num_clients: int, num_samples: Optional[Sequence[float]] = None) -> np.ndarray:
    """Per-client scale factors handed to the aggregation kernel (sum to 1)."""
    if num_samples is None:
        return np.full((num_clients,), 1.0 / num_clients, np.float32)
    ns = np.asarray(num_samples, np.float64)
    if ns.shape != (num_clients,) or ns.sum() <= 0:
        raise ValueError("bad sample counts")
    return (ns / ns.sum()).astype(np.float32)


@dataclass
class FedAvg:
    """w <- sum_k c_k w_k.  ``weighted=False`` reproduces the reference (uniform 1/N)."""

    weighted: bool = False
    name: str = "fedavg"

    def aggregate(self, global_w: WeightList, updates: Sequence[WeightList],
                  num_samples: Optional[Sequence[float]] = None) -> WeightList:
        if self.weighted and num_samples is not None:
            return fedavg_weighted(updates, num_samples)
        return fedavg_mean(updates)

    def state(self) -> dict:
        return {}

    def load_state(self, st: dict):
        pass


@dataclass
class FedAdam:
    """Adaptive server optimiser (Reddi et al., "Adaptive Federated Optimization").

    delta = sum_k c_k w_k - w;  m = b1 m + (1-b1) delta;  v = b2 v + (1-b2) delta^2;
    w <- w + eta * m / (sqrt(v) + tau).  No bias correction (as in the paper).
    Non-trainable tensors (BN moving statistics) are plain-averaged: ``trainable_mask``.
    """

    eta: float = 1e-2
    beta1: float = 0.9
    beta2: float = 0.99
    tau: float = 1e-3
    weighted: bool = True
    name: str = "fedadam"
    trainable_mask: Optional[List[bool]] = None
    m: Optional[WeightList] = field(default=None, repr=False)
    v: Optional[WeightList] = field(default=None, repr=False)

    def aggregate(self, global_w: WeightList, updates: Sequence[WeightList],
                  num_samples: Optional[Sequence[float]] = None) -> WeightList:
        avg = (fedavg_weighted(updates, num_samples) if (self.weighted and num_samples is not None)
               else fedavg_mean(updates))
        if self.m is None:
            self.m = [np.zeros_like(a) for a in avg]
            self.v = [np.zeros_like(a) for a in avg]
        out = []
        for i, (w, a) in enumerate(zip(global_w, avg)):
            if self.trainable_mask is not None and not self.trainable_mask[i]:
                out.append(a)
                continue
            w = np.asarray(w, np.float32)
            d = a - w
            self.m[i] = np.float32(self.beta1) * self.m[i] + np.float32(1 - self.beta1) * d
            self.v[i] = np.float32(self.beta2) * self.v[i] + np.float32(1 - self.beta2) * d * d
            out.append(w + np.float32(self.eta) * self.m[i] / (np.sqrt(self.v[i]) + np.float32(self.tau)))
        return out

    def state(self) -> dict:
        return {"m": self.m, "v": self.v}

    def load_state(self, st: dict):
        self.m, self.v = st.get("m"), st.get("v")


def make_strategy(name: str, **kw):
    name = name.lower()
    if name in ("fedavg", "mean"):
        return FedAvg(weighted=kw.pop("weighted", False))
    if name in ("fedavg_weighted", "weighted"):
        return FedAvg(weighted=True)
    if name == "fedadam":
        return FedAdam(**kw)
    raise ValueError(f"unknown aggregation strategy {name!r}")
