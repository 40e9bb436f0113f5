"""Observability: JSONL records, TensorBoard event files, CUDA-event phase timers.

The reference logs with ``print`` and one Keras callback,
``TensorBoard(log_dir="send_logs/logs/<ts>-<round>", histogram_freq=1)``
(``client_fit_model.py:153-154,166``): per-epoch scalars (loss, accuracy, val_loss,
val_accuracy) and per-layer weight histograms.  ``RoundLogger`` writes the same event
files (through ``torch.utils.tensorboard``, no TensorFlow needed) plus a structured
JSONL stream.
"""
from __future__ import annotations

import json
import os
import time
from contextlib import contextmanager
from datetime import datetime
from typing import Dict, List, Optional, Sequence

import numpy as np


class JsonlWriter:
    def __init__(self, path: Optional[str]):
        self.path = path
        if path:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)

    def write(self, **rec):
        if not self.path:
            return
        rec.setdefault("ts", time.time())
        with open(self.path, "a") as f:
            f.write(json.dumps(rec, default=float) + "\n")


class RoundLogger:
    """Per-round TensorBoard + JSONL sink, directory layout as the reference's callback."""

    def __init__(self, log_root: str, round_idx: int, tensorboard: bool = True, histogram_freq: int = 1,
                 tensor_names: Optional[Sequence[str]] = None):
        stamp = datetime.now().strftime("%Y%m%d-%H%M%S")
        self.dir = os.path.join(log_root, f"{stamp}-{round_idx}")  # client_fit_model.py:153
        self.histogram_freq = histogram_freq
        self.tensor_names = list(tensor_names or [])
        self.jsonl = JsonlWriter(os.path.join(self.dir, "metrics.jsonl"))
        self.tb_train = self.tb_val = None
        if tensorboard:
            try:
                from torch.utils.tensorboard import SummaryWriter
                self.tb_train = SummaryWriter(os.path.join(self.dir, "train"))
                self.tb_val = SummaryWriter(os.path.join(self.dir, "validation"))
            except Exception as e:  # tensorboard missing -> JSONL only
                self.jsonl.write(event="tensorboard_unavailable", error=repr(e))

    def epoch(self, epoch: int, train: Dict[str, float], val: Optional[Dict[str, float]] = None,
              weights: Optional[Sequence[np.ndarray]] = None):
        self.jsonl.write(event="epoch", epoch=epoch, **{f"train_{k}": v for k, v in train.items()},
                         **({f"val_{k}": v for k, v in val.items()} if val else {}))
        if self.tb_train:
            for k, v in train.items():
                self.tb_train.add_scalar(f"epoch_{k}", v, epoch)
            if val:
                for k, v in val.items():
                    self.tb_val.add_scalar(f"epoch_{k}", v, epoch)
            if weights is not None and self.histogram_freq and epoch % self.histogram_freq == 0:
                for name, w in zip(self.tensor_names, weights):
                    self.tb_train.add_histogram(name.replace(":0", "_0"), np.asarray(w).reshape(-1), epoch)

    def close(self):
        for tb in (self.tb_train, self.tb_val):
            if tb:
                tb.flush()
                tb.close()


class PhaseTimer:
    """Wall-clock (CPU) or CUDA-event (GPU) timing of named phases; ``summary()`` gives ms."""

    def __init__(self, cuda: bool = False):
        self.cuda = cuda
        self.ms: Dict[str, float] = {}
        self._pending: List[tuple] = []

    @contextmanager
    def phase(self, name: str):
        if self.cuda:
            import torch
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            yield
            b.record()
            self._pending.append((name, a, b))
        else:
            t0 = time.perf_counter()
            yield
            self.ms[name] = self.ms.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def summary(self) -> Dict[str, float]:
        if self.cuda and self._pending:
            import torch
            torch.cuda.synchronize()
            for name, a, b in self._pending:
                self.ms[name] = self.ms.get(name, 0.0) + a.elapsed_time(b)
            self._pending.clear()
        return dict(self.ms)
