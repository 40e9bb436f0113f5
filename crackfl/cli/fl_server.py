"""``python -m crackfl.cli.fl_server`` - the FedAvg coordinator (reference entry point: ``python fl_server.py``,
``fl_server.py:229-232``).  Builds the global model (the same U-Net the clients train, not the leftover
``mobilenet_v2`` arguments of ``fl_server.py:230-231``), then serves the gRPC protocol until FIN."""
from __future__ import annotations

import argparse
import sys

from ..protocol.server import FLServer, serve_forever
from ..trainer import spec_from_config
from .common import parse_with_config


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--weights", default=None, help="initial weights: Keras .h5, weights.pickle or checkpoint dir")
    args, cfg = parse_with_config(ap, argv)
    spec = spec_from_config(cfg)
    weights = load_any_weights(args.weights, spec, cfg) if args.weights else spec.init_weights(cfg.server.seed)
    aggregator = None
    kind = cfg.server.aggregator
    if kind == "auto":
        import torch
        kind = "device" if torch.cuda.is_available() else "host"
    if kind == "device":
        from ..parallel.device_agg import DeviceAggregator
        aggregator = DeviceAggregator(spec, cfg.server.strategy, eta=cfg.server.fedadam_eta, beta1=cfg.server.fedadam_beta1,
                                      beta2=cfg.server.fedadam_beta2, tau=cfg.server.fedadam_tau)
    server = FLServer(cfg, weights, spec, aggregator=aggregator)
    serve_forever(server)
    return 0


def load_any_weights(path: str, spec, cfg=None):
    import os
    from ..io import checkpoint as ckpt
    from ..io import weights as wio
    if os.path.isdir(path):
        return ckpt.load_checkpoint(path, spec)["weights"]
    with open(path, "rb") as f:
        head = f.read(8)
    if head == b"\x89HDF\r\n\x1a\n":
        from ..io.keras_h5 import load_keras_h5
        return load_keras_h5(path, spec)
    with open(path, "rb") as f:
        blob = f.read()
    if head == wio.RAW_MAGIC:
        return wio.loads_raw(blob)
    return wio.loads_pickle(blob)  # restricted unpickler


if __name__ == "__main__":
    sys.exit(main())
