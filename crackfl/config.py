"""Configuration: every literal of the reference exposed as a knob.

Defaults are the reference's *effective* values; the citation beside each field is where
the literal lives in the reference.  Configs load from YAML/JSON files, from
``CRACKFL_*`` environment variables and from ``--section.key=value`` CLI overrides.
"""
from __future__ import annotations

import dataclasses
import json
import os
from dataclasses import dataclass, field, fields, is_dataclass
from typing import Any, Dict, List, Optional, Tuple


@dataclass
class TransportConfig:
    bind: str = "[::]:8889"  # fl_server.py:218
    target: str = "localhost:8889"  # fl_client.py:181
    max_message_mib: int = 512  # fl_server.py:215 (send option was misspelt there, quirk Q12; set both here)
    workers: int = 10  # fl_server.py:216
    service_path: str = "transport.TransportService"  # package name of the lost .proto is unknown (H8)
    payload_format: str = "raw"  # "raw" (safe default) | "pickle" (reference wire format, opt-in)
    allow_pickle: bool = False  # accept pickle payloads from peers (RCE surface: fl_server.py:179)
    chunk_mib: int = 100  # fl_client.py:36 (log files); also used to chunk weights > max_message


@dataclass
class ServerConfig:
    max_rounds: int = 5  # fl_server.py:18 MAX_NUM_ROUND
    admission_window_s: float = 10.0  # fl_server.py:42
    ready_delay_s: float = 5.0  # fl_server.py:56
    initial_model_version: int = 1  # fl_server.py:17
    model_type: str = "unet"  # reference says "mobilenet_v2" (fl_server.py:75, quirk Q14); alias accepted
    strategy: str = "fedavg"  # fedavg | fedavg_weighted | fedadam (last two are extensions)
    fedadam_eta: float = 1e-2
    fedadam_beta1: float = 0.9
    fedadam_beta2: float = 0.99
    fedadam_tau: float = 1e-3
    min_fit_clients: int = 0  # extension: quorum for round_timeout_s (0 = all registered, as the reference)
    round_timeout_s: float = 0.0  # extension: 0 = wait forever, as the reference (fl_server.py:116)
    weights_dir: str = "./server_weights"  # fl_server.py:104
    log_dir: str = "."  # 'L' uploads land in '.' + title[11:] (fl_server.py:84-89)
    resume: bool = False  # extension: reload round/version/weights/optimizer state
    aggregator: str = "auto"  # auto | host | device
    check_finite: bool = True  # extension: reject updates holding NaN/Inf before they are queued
    fin_grace_s: float = 0.0  # serve_forever: how long to keep answering after the last round; 0 = client.poll_interval_s + 5
    seed: int = 0


@dataclass
class ClientConfig:
    poll_interval_s: float = 20.0  # fl_client.py:141 (comment there says 30)
    weights_path: str = "./saved_weight/weights.pickle"  # client_fit_model.py:239
    write_weights_file: bool = True  # the reference hands weights over through this file (fl_client.py:124)
    name: str = ""  # "" -> client{randint(1,100000)} as fl_client.py:26
    upload_logs: bool = False  # fl_client.py:110-118 is commented out in the reference
    log_root: str = "./send_logs/logs"  # client_fit_model.py:153


@dataclass
class TrainConfig:
    image_size: Tuple[int, int] = (128, 128)  # client_fit_model.py:55
    batch_size: int = 16  # client_fit_model.py:56 (ctor arg bs=128 is ignored there, quirk Q6)
    local_epochs: int = 10  # client_fit_model.py:166 (ctor arg ec=1 ignored)
    lr: float = 1e-3  # Keras Adam default
    beta1: float = 0.9
    beta2: float = 0.999
    adam_eps: float = 1e-7
    loss: str = "bce"  # "bce" (client_fit_model.py:157) | "bce_dice" (extension)
    bce_clip: bool = False  # False: cross-entropy from the logit, what Keras 2 fit() computes for a sigmoid output layer
    #                         True: clipped probabilities [1e-7, 1-1e-7] (backend fall-back / Keras 3): no gradient outside
    dtype: str = "bf16"  # compute dtype of the B200 engine; the oracle back-end is fp32
    backend: str = "auto"  # auto | b200 | torch_ref
    shuffle_batches: bool = True  # Keras fit on a Sequence shuffles batch order each epoch
    validate: bool = True  # validation_data=val_gen every epoch
    predict_round: int = 5  # client_fit_model.py:235
    tensorboard: bool = True  # client_fit_model.py:153-154
    histogram_freq: int = 1
    width_mult: int = 1  # extension: "wide U-Net"
    cuda_graph: bool = True
    seed: int = 1337


@dataclass
class DataConfig:
    train_image_dir: str = ""  # client_fit_model.py:58 (absolute path on the author's machine)
    train_mask_dir: str = ""  # client_fit_model.py:59
    predict_image_dir: str = ""  # client_fit_model.py:183
    split_seed: int = 1337  # client_fit_model.py:76
    split_count: int = 6213  # client_fit_model.py:76-82 ("val_samples", used as the TRAIN count there)
    split_mode: str = "client"  # "client": train=first N (client_fit_model.py:78-81); "script": val=last N (Segmentation.py:84-90)
    synthetic: bool = True  # dataset is not shipped; synthetic crack images are the only input
    synthetic_train: int = 6213
    synthetic_val: int = 1024
    workers: int = 4
    prefetch_batches: int = 16  # extension: batches decoded ahead of the training loop (0 = inline, as the reference)
    prefetch_threads: int = 8  # extension: threads that build those batches (whole batches in parallel)


@dataclass
class MeasureConfig:
    threshold: int = 127  # test/Segmentation2.py:119
    eps_fine: float = 0.01  # test/Segmentation2.py:127
    eps_coarse: float = 0.1  # test/Segmentation2.py:128
    out_dir: str = "contour"  # test/Segmentation2.py:139-141
    write_images: bool = True


@dataclass
class Config:
    transport: TransportConfig = field(default_factory=TransportConfig)
    server: ServerConfig = field(default_factory=ServerConfig)
    client: ClientConfig = field(default_factory=ClientConfig)
    train: TrainConfig = field(default_factory=TrainConfig)
    data: DataConfig = field(default_factory=DataConfig)
    measure: MeasureConfig = field(default_factory=MeasureConfig)

    # ---------------------------------------------------------------------------------
    def to_dict(self) -> Dict[str, Any]:
        return dataclasses.asdict(self)

    def update(self, d: Dict[str, Any]) -> "Config":
        for sec, vals in d.items():
            if not hasattr(self, sec):
                raise KeyError(f"unknown config section {sec!r}")
            obj = getattr(self, sec)
            for k, v in vals.items():
                _set(obj, k, v)
        return self

    def override(self, dotted: str, value: str) -> "Config":
        sec, _, key = dotted.partition(".")
        if not key or not hasattr(self, sec):
            raise KeyError(f"bad override {dotted!r} (want section.key)")
        _set(getattr(self, sec), key, value)
        return self

    @classmethod
    def load(cls, path: Optional[str] = None, overrides: Optional[List[str]] = None, env: bool = True) -> "Config":
        cfg = cls()
        if path:
            with open(path) as f:
                txt = f.read()
            if path.endswith((".yml", ".yaml")):
                import yaml
                cfg.update(yaml.safe_load(txt) or {})
            else:
                cfg.update(json.loads(txt))
        if env:
            for k, v in os.environ.items():
                if k.startswith("CRACKFL_") and "__" in k:
                    sec, key = k[len("CRACKFL_"):].lower().split("__", 1)
                    if hasattr(cfg, sec) and hasattr(getattr(cfg, sec), key):
                        _set(getattr(cfg, sec), key, v)
        for ov in overrides or []:
            k, _, v = ov.lstrip("-").partition("=")
            cfg.override(k, v)
        return cfg


def _coerce(cur: Any, v: Any) -> Any:
    if isinstance(v, str) and not isinstance(cur, str):
        if isinstance(cur, bool):
            return v.strip().lower() in ("1", "true", "yes", "on")
        if isinstance(cur, int):
            return int(v)
        if isinstance(cur, float):
            return float(v)
        if isinstance(cur, (tuple, list)):
            parts = [p for p in v.replace("x", ",").split(",") if p]
            return tuple(int(p) for p in parts)
    if isinstance(cur, tuple) and isinstance(v, list):
        return tuple(v)
    return v


def _set(obj: Any, key: str, v: Any):
    if not is_dataclass(obj) or key not in {f.name for f in fields(obj)}:
        raise KeyError(f"unknown config key {type(obj).__name__}.{key}")
    setattr(obj, key, _coerce(getattr(obj, key), v))


def fast_test_config() -> Config:
    """Reference protocol with the 5/10/20 s sleeps removed (tests, benchmarks)."""
    c = Config()
    c.server.ready_delay_s = 0.0
    c.server.admission_window_s = 0.5
    c.client.poll_interval_s = 0.01
    return c
