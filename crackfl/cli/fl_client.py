"""``python -m crackfl.cli.fl_client`` - one federated client (reference: ``python fl_client.py``,
``fl_client.py:178-188``): connects, registers, fits locally each round, uploads its update."""
from __future__ import annotations

import argparse
import json
import sys

from ..protocol.client import FLClient, make_stub
from ..trainer import LocalTrainer
from .common import parse_with_config


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--client-index", type=int, default=0, help="selects the synthetic shard / RNG stream")
    args, cfg = parse_with_config(ap, argv)
    predict_ds = None
    if cfg.data.predict_image_dir:
        from ..data.generator import Generator, list_jpgs
        paths = list_jpgs(cfg.data.predict_image_dir)
        predict_ds = Generator(paths, paths, cfg.train.batch_size, cfg.train.image_size)
    trainer = LocalTrainer(cfg, predict_ds=predict_ds, client_index=args.client_index)
    call, channel = make_stub(cfg)
    client = FLClient(cfg, trainer, call)
    summary = client.run()
    channel.close()
    print(json.dumps({k: v for k, v in summary.items() if k != "history"}))
    return 0 if summary.get("state") != "CTW" else 2


if __name__ == "__main__":
    sys.exit(main())
