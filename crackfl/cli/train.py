"""``python -m crackfl.cli.train`` - centralised training (reference: ``python test/Segmentation.py``):
same data split / model / Adam+BCE, ``ModelCheckpoint("crack_segmentation.h5", save_best_only=True)`` on val_loss
(``test/Segmentation.py:177-186``, 60 epochs) and a final native checkpoint in place of the TF SavedModel directory
(``model.save("my_model")``, ``:187`` - needs TensorFlow, out of scope)."""
from __future__ import annotations

import argparse
import json
import random
import sys

from ..io import checkpoint as ckpt
from ..io.keras_h5 import save_keras_h5
from ..trainer import make_backend, spec_from_config
from ..utils.metrics import RoundLogger
from .common import parse_with_config


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--epochs", type=int, default=60)  # test/Segmentation.py:185
    ap.add_argument("--out", default="crack_segmentation.h5")
    ap.add_argument("--save-dir", default="my_model")
    ap.add_argument("--init", default=None)
    args, cfg = parse_with_config(ap, argv)
    cfg.data.split_mode = "script"  # test/Segmentation.py:84-90: val = LAST 6213
    spec = spec_from_config(cfg)
    from ..data.generator import make_datasets
    train_ds, val_ds = make_datasets(cfg)
    be = make_backend(cfg, spec)
    from .fl_server import load_any_weights
    be.begin_round(load_any_weights(args.init, spec) if args.init else spec.init_weights(cfg.server.seed))
    logger = RoundLogger("logs", 0, cfg.train.tensorboard, 0, spec.names())
    rng = random.Random(cfg.train.seed)
    order = list(range(len(train_ds)))
    best = float("inf")
    for epoch in range(args.epochs):
        rng.shuffle(order)
        be.reset_metrics()
        for bi in order:
            be.train_batch(*train_ds[bi])
        tm = be.read_metrics()
        be.reset_metrics()
        for bi in range(len(val_ds)):
            be.eval_batch(*val_ds[bi])
        vm = be.read_metrics()
        logger.epoch(epoch, tm, vm)
        improved = vm["loss"] < best
        if improved:  # ModelCheckpoint(save_best_only=True) monitors val_loss
            best = vm["loss"]
            save_keras_h5(args.out, spec, be.get_weights(), full_model=True, input_hw=cfg.train.image_size)
        print(f"Epoch {epoch + 1}/{args.epochs} - loss {tm['loss']:.4f} - accuracy {tm['accuracy']:.4f} - val_loss "
              f"{vm['loss']:.4f} - val_accuracy {vm['accuracy']:.4f}" + (f"  (saved {args.out})" if improved else ""))
    ckpt.save_checkpoint(args.save_dir, spec, be.get_weights(), meta={"epochs": args.epochs, "best_val_loss": best})
    logger.close()
    print(json.dumps({"best_val_loss": best, "h5": args.out, "checkpoint": args.save_dir}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
