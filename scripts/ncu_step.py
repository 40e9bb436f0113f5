#!/usr/bin/env python
"""Target program for `ncu --set full`: two eager train steps of the 256x256, B=16 U-Net on one stream (every kernel of
the step once per step), then the single-GPU forms of the aggregation kernel and the loss/measure/synth kernels.
Run plain first (must exit 0), then under ncu with `-s <launches of step 1>` so that only the second step is captured."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.data.synthetic_device import synth_crack_batch_device
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.models.unet import UNetEngine
from crackfl.parallel.device_agg import DeviceAggregator
from crackfl.postproc.device_measure import measure_masks_device

B, H = int(os.environ.get("B", 16)), int(os.environ.get("SIZE", 256))
os.environ.setdefault("CRACKFL_SIDE_STREAM", "0")
eng = UNetEngine(S, B, H, H, cuda_graph=False)
eng.load_weights(S.init_weights(0)); eng.reset_optimizer()
x, y = synth_crack_batch_device(B, H, H, 1)
eng.set_batch(x, y)
for _ in range(int(os.environ.get("STEPS", 2))):
    eng.train_step()
torch.cuda.synchronize()
if os.environ.get("EXTRAS", "1") != "0":
    eng.eval_step(); p = eng.predict()
    measure_masks_device(p)
    import numpy as np
    agg = DeviceAggregator(S, strategy="fedadam", payload="fp32", max_clients=8)
    w = S.init_weights(0)
    agg.aggregate(w, [[a + 0.01 for a in w] for _ in range(8)], [1.0] * 8)
    torch.cuda.synchronize()
print("launches_per_step", eng.launches_per_step)
