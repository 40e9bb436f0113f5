"""Smallest end-to-end exercise of every kernel, for compute-sanitizer (memcheck / racecheck / synccheck)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from crackfl.data.synthetic_device import synth_crack_batch_device
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.models.unet import UNetEngine
from crackfl.parallel.device_agg import DeviceAggregator
from crackfl.postproc.device_measure import measure_masks_device

eng = UNetEngine(S, 2, 64, 64, cuda_graph=False)
eng.load_weights(S.init_weights(0)); eng.reset_optimizer()
x, y = synth_crack_batch_device(2, 64, 64, 3)
for _ in range(2):
    eng.reset_metrics(); eng.train_step(x, y)
eng.eval_step(x, y)
p = eng.predict(x)
m = measure_masks_device(p)
torch.cuda.synchronize()
w = eng.get_weights()
agg = DeviceAggregator(S, "fedadam", payload="bf16_delta", max_clients=2)
out = agg.aggregate(w, [w, [a + 0.01 for a in w]], [1, 3])
print("SANITIZE_RUN_OK", eng.read_metrics(), float(m.sum()), float(out[0].mean()))
