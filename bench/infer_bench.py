#!/usr/bin/env python
"""Inference throughput of the engine (predict = forward in eval mode + sigmoid probabilities at full resolution, the
path behind `learning_fit.Predict`, client_fit_model.py:176-223) and of predict + device-side mask measurement.
CUDA events, warm-up, inputs rotate over 8 device-resident uint8 batches.  One JSON line."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.data.synthetic_device import synth_crack_batch_device
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.models.unet import UNetEngine

ap = argparse.ArgumentParser()
ap.add_argument("--size", type=int, default=256)
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--steps", type=int, default=300)
ap.add_argument("--warmup", type=int, default=20)
a = ap.parse_args()
B, H = a.batch, a.size
eng = UNetEngine(S, B, H, H)
eng.load_weights(S.init_weights(0))
xs = [synth_crack_batch_device(B, H, H, s)[0] for s in range(8)]


def timed(fn):
    for i in range(a.warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.steps


ms_pred = timed(lambda i: eng.predict(xs[i % 8]))
out = {"metric": "U-Net predict images/s (eval-mode forward + probabilities)", "batch": B, "image": f"{H}x{H}",
       "ms_per_batch": ms_pred, "images_per_s": B / ms_pred * 1e3, "dtype": "bf16"}
try:
    from crackfl.postproc.device_measure import measure_masks_device
    ms_meas = timed(lambda i: measure_masks_device(eng.predict(xs[i % 8])))
    out["ms_per_batch_with_device_measure"] = ms_meas
    out["images_per_s_with_device_measure"] = B / ms_meas * 1e3
except Exception as e:  # pragma: no cover
    out["device_measure_error"] = repr(e)
print(json.dumps(out))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/infer_bench.json", "w"), indent=1)
