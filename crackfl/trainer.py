"""Local trainer: the ``learning_fit`` class of ``client_fit_model.py:46-240``.

``LocalTrainer.fit(weights, round)`` is ``manage_train`` + ``train_model_tosave``
(``client_fit_model.py:152-174,225-240``): every round starts from the incoming global
weights with FRESH Adam state (the reference rebuilds model+optimizer per round, quirk
Q5; here only the moments are zeroed), runs ``local_epochs`` epochs over the training
``Sequence`` in shuffled batch order, validates every epoch, logs TensorBoard scalars +
weight histograms, runs predict+measurement when ``round == predict_round``
(``client_fit_model.py:235``) and returns the weights (the reference returns ``None``
and hands them over through a pickle file, quirk Q8).

Back-ends:

* ``b200``      - ``crackfl.models.unet.UNetEngine``: hand-written sm_100a kernels, bf16
                  activations, CUDA-graphed train step.  Default whenever a GPU is visible;
                  raises if the in-tree CUDA library is missing (no silent fallback).
* ``torch_ref`` - the fp32 PyTorch oracle (``models/unet_ref.py``); CPU plumbing config only.
"""
from __future__ import annotations

import random
import time
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .config import Config
from .io import weights as wio
from .models.spec import REFERENCE_SPEC, UNetSpec, wide_spec
from .utils.metrics import RoundLogger


def spec_from_config(cfg: Config) -> UNetSpec:
    return REFERENCE_SPEC if cfg.train.width_mult == 1 else wide_spec(cfg.train.width_mult)


class TorchRefBackend:
    """fp32 oracle back-end (CPU)."""

    name = "torch_ref"

    def __init__(self, cfg: Config, spec: UNetSpec):
        self.cfg, self.spec, self.tr = cfg, spec, None
        self._acc = [0.0, 0.0, 0]

    def begin_round(self, weights: Optional[Sequence[np.ndarray]]):
        from .models.unet_ref import RefTrainer
        if weights is None:  # "the model is already where it is trained": carry on from the current weights
            weights = self.tr.get_weights()
        self.tr = RefTrainer(weights, self.spec, lr=self.cfg.train.lr, loss=self.cfg.train.loss,
                             bce_clip=self.cfg.train.bce_clip)
        self.reset_metrics()

    @staticmethod
    def _prep(x_u8: np.ndarray, y_u8: Optional[np.ndarray]):
        import torch
        x = torch.from_numpy(np.ascontiguousarray(x_u8)).permute(0, 3, 1, 2).to(torch.float32) / 255.0
        y = None if y_u8 is None else torch.from_numpy(np.ascontiguousarray(y_u8)).permute(0, 3, 1, 2).to(torch.float32)
        return x, y

    def reset_metrics(self):
        self._acc = [0.0, 0.0, 0]

    def train_batch(self, x_u8, y_u8):
        l, a = self.tr.train_step(*self._prep(x_u8, y_u8))
        self._acc[0] += l; self._acc[1] += a; self._acc[2] += 1

    def eval_batch(self, x_u8, y_u8):
        l, a = self.tr.eval_step(*self._prep(x_u8, y_u8))
        self._acc[0] += l; self._acc[1] += a; self._acc[2] += 1

    def read_metrics(self) -> Dict[str, float]:
        n = max(self._acc[2], 1)
        return {"loss": self._acc[0] / n, "accuracy": self._acc[1] / n}

    def predict_batch(self, x_u8) -> np.ndarray:
        x, _ = self._prep(x_u8, None)
        return self.tr.predict(x).permute(0, 2, 3, 1).numpy()

    def get_weights(self) -> List[np.ndarray]:
        return self.tr.get_weights()


def make_backend(cfg: Config, spec: UNetSpec, device: Optional[int] = None, arena=None):
    kind = cfg.train.backend
    if kind == "auto":
        import torch
        kind = "b200" if torch.cuda.is_available() else "torch_ref"
    if kind == "torch_ref":
        return TorchRefBackend(cfg, spec)
    if kind == "b200":
        from .models.unet import B200Backend  # raises loudly if the CUDA library is missing
        return B200Backend(cfg, spec, device=device, arena=arena)
    raise ValueError(f"unknown backend {kind!r}")


def _prefetched(ds, order, depth: int, threads: int = 1):
    """Batches of ``ds`` in ``order``, decoded up to ``depth`` batches ahead by ``threads`` worker threads.

    The reference decodes the 32 JPEGs of a step inside ``fit``'s loop (``client_fit_model.py:30-43`` is called by Keras
    with no ``workers``), so decode + resize sit on the critical path of every step.  Here every worker takes the next
    batch index, builds that whole batch (cv2 releases the GIL while it decodes) and parks it in an ordered buffer; the
    training loop only ever waits for the next batch in order.  Parallelism is per BATCH, not per image: one Python call
    per batch and no nested pools.  Device-resident datasets and ``depth <= 0`` fall through to plain indexing."""
    if depth <= 0 or len(order) < 2 or getattr(ds, "device_resident", False):
        for bi in order:
            yield ds[bi]
        return
    import threading
    cv = threading.Condition()
    results, state = {}, {"next": 0, "consumed": 0, "stop": False}
    n = len(order)

    def work():
        while True:
            with cv:
                while not state["stop"] and state["next"] < n and state["next"] - state["consumed"] >= depth:
                    cv.wait()
                if state["stop"] or state["next"] >= n:
                    return
                i = state["next"]
                state["next"] += 1
            try:
                item = ("ok", ds[order[i]])
            except BaseException as e:  # surfaced in the consumer, in order
                item = ("err", e)
            with cv:
                results[i] = item
                cv.notify_all()

    ths = [threading.Thread(target=work, daemon=True, name=f"crackfl-prefetch-{k}") for k in range(max(1, threads))]
    for t in ths:
        t.start()
    try:
        for i in range(n):
            with cv:
                while i not in results:
                    cv.wait()
                kind, item = results.pop(i)
                state["consumed"] = i + 1
                cv.notify_all()
            if kind == "err":
                raise item
            yield item
    finally:
        with cv:
            state["stop"] = True
            cv.notify_all()
        for t in ths:
            t.join(1.0)


class LocalTrainer:
    """``learning_fit(mt, ec, bs, pr, cr)``; the ignored ctor args of the reference (quirk Q6)
    are real configuration here (``cfg.train.local_epochs`` / ``batch_size``)."""

    def __init__(self, cfg: Config, train_ds=None, val_ds=None, predict_ds=None, spec: Optional[UNetSpec] = None,
                 backend=None, client_index: int = 0, log=print):
        self.cfg = cfg
        self.spec = spec or spec_from_config(cfg)
        if train_ds is None:
            from .data.generator import make_datasets
            train_ds, val_ds = make_datasets(cfg, client_index)
        self.train_ds, self.val_ds, self.predict_ds = train_ds, val_ds, predict_ds
        self.backend = backend or make_backend(cfg, self.spec)
        self.log = log
        self.rng = random.Random(cfg.train.seed + client_index)
        self.last_measurements = None

    # reference-compatible entry point: bytes in, bytes out
    def manage_train(self, params: Optional[bytes] = None, cr: Optional[int] = None, initial: Optional[bytes] = None):
        blob = params if params is not None else initial
        if blob is None:
            return []
        w, _ = self.fit(wio.loads_auto(blob, allow_pickle=self.cfg.transport.allow_pickle), cr or 0)
        return w

    def fit(self, weights: Optional[Sequence[np.ndarray]], round_idx: int,
            return_weights: bool = True) -> Tuple[Optional[List[np.ndarray]], Dict[str, float]]:
        """``weights=None``: the round's global model is already resident in the back-end (the on-box runtime's
        broadcast lands in the engine's arena), only the optimiser and the metrics are reset.  ``return_weights=False``
        skips the device-to-host export for callers that read the arena in place."""
        t = self.cfg.train
        be = self.backend
        self.log(f"### Model Training - Round: {round_idx} ({be.name}) ###")
        be.begin_round(weights)
        logger = RoundLogger(self.cfg.client.log_root, round_idx, t.tensorboard, t.histogram_freq, self.spec.names()) \
            if (t.tensorboard or self.cfg.client.log_root) else None
        n_batches = len(self.train_ds)
        order = list(range(n_batches))
        t0 = time.perf_counter()
        train_m, val_m = {}, None
        for epoch in range(t.local_epochs):
            if t.shuffle_batches:
                self.rng.shuffle(order)  # Keras shuffles batch ORDER of a Sequence, not contents
            be.reset_metrics()
            for x, y in _prefetched(self.train_ds, order, self.cfg.data.prefetch_batches, self.cfg.data.prefetch_threads):
                be.train_batch(x, y)
            train_m = be.read_metrics()
            if t.validate and self.val_ds is not None and len(self.val_ds) > 0:
                be.reset_metrics()
                for bi in range(len(self.val_ds)):
                    x, y = self.val_ds[bi]
                    be.eval_batch(x, y)
                val_m = be.read_metrics()
            if logger:
                logger.epoch(epoch, train_m, val_m, be.get_weights() if t.histogram_freq else None)
            self.log(f"epoch {epoch + 1}/{t.local_epochs} - loss {train_m['loss']:.4f} - accuracy {train_m['accuracy']:.4f}"
                     + (f" - val_loss {val_m['loss']:.4f} - val_accuracy {val_m['accuracy']:.4f}" if val_m else ""))
        if hasattr(be, "finish"):
            be.finish()  # device back-end: the fit time below must cover the queued GPU work
        fit_s = time.perf_counter() - t0
        out = be.get_weights() if return_weights else None
        if round_idx == t.predict_round and self.predict_ds is not None:
            from .postproc.measure import predict_and_measure
            self.last_measurements = predict_and_measure(be.predict_batch, self.predict_ds, self.cfg.measure)
            self.log(" predict done ")
        if logger:
            logger.close()
        n_img = n_batches * t.batch_size * t.local_epochs
        metrics = {"loss": train_m.get("loss"), "accuracy": train_m.get("accuracy"),
                   "val_loss": val_m["loss"] if val_m else None, "val_accuracy": val_m["accuracy"] if val_m else None,
                   "num_samples": n_batches * t.batch_size, "fit_seconds": fit_s,
                   "images_per_sec": n_img / fit_s if fit_s > 0 else 0.0}
        return out, metrics
