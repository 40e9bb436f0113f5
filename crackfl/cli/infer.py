"""``python -m crackfl.cli.infer`` - inference + crack measurement (reference: ``python test/Segmentation2.py``):
load a model, predict, write ``pred*.png``, threshold/contour/measure, write ``contour/img{1,2,3}.jpg``
(``test/Segmentation2.py:94-144``)."""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

from ..postproc.measure import measure, prob_to_bgr
from ..trainer import make_backend, spec_from_config
from .common import parse_with_config


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--weights", required=True, help="Keras .h5, weights.pickle or checkpoint directory")
    ap.add_argument("--images", default=None, help="directory of .jpg images (default: synthetic validation batch)")
    ap.add_argument("--index", type=int, default=1, help="which prediction to measure (reference uses val_preds[1])")
    ap.add_argument("--out-dir", default=".")
    args, cfg = parse_with_config(ap, argv)
    spec = spec_from_config(cfg)
    from .fl_server import load_any_weights
    weights = load_any_weights(args.weights, spec)
    be = make_backend(cfg, spec)
    be.begin_round(weights)
    if args.images:
        from ..data.generator import Generator, list_jpgs
        paths = list_jpgs(args.images)
        ds = Generator(paths, paths, cfg.train.batch_size, cfg.train.image_size)
        x = ds.images_only(0) if len(ds) else np.stack([ds._load_x(p) for p in paths])
    else:
        from ..data.generator import synth_crack_batch
        x, _ = synth_crack_batch(cfg.train.batch_size, cfg.train.image_size[0], cfg.train.image_size[1], 4242)
    if len(x) < cfg.train.batch_size:
        x = np.concatenate([x, np.repeat(x[-1:], cfg.train.batch_size - len(x), 0)])
    preds = be.predict_batch(x)
    import cv2
    os.makedirs(args.out_dir, exist_ok=True)
    rows = []
    for i, p in enumerate(preds):
        bgr = prob_to_bgr(p)
        if i == args.index:
            cv2.imwrite(os.path.join(args.out_dir, "pred2.png"), bgr)  # test/Segmentation2.py:105
        m = measure(bgr, cfg.measure.threshold, cfg.measure.eps_fine, cfg.measure.eps_coarse,
                    os.path.join(args.out_dir, cfg.measure.out_dir), write_images=(i == args.index))
        rows.append({"index": i, "area": m.area, "length": m.length, "round2": m.round2, "contours": m.mount})
    sel = rows[min(args.index, len(rows) - 1)]
    print(f" contour area : {sel['area']}\n contour length : {sel['length']}")
    print(json.dumps({"measurements": rows}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
