#!/usr/bin/env python
"""Memory-roofline estimate of one train step: for every launch of the schedule, the bytes of the distinct activation
buffers it is handed (each counted once per launch, whole buffer), summed and divided by the measured HBM copy bandwidth
(MEASURED_PEAKS.json).  Weights / gradients / statistics (a few MB in total) are ignored.  Needs a GPU (the engine
allocates its buffers on construction).  Writes gpurun_out/step_bytes.json."""
import collections, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.models.unet import UNetEngine

B, H = int(os.environ.get("B", 16)), int(os.environ.get("SIZE", 256))
eng = UNetEngine(S, B, H, H, cuda_graph=False)
skip = {"ops", "prep_descs", "wgrad_ws", "wgrad_descs", "zero_region"}
bufs = {t.data_ptr(): (n, t.numel() * t.element_size()) for n, t in eng._bufs.items() if n not in skip}
per_kernel = collections.defaultdict(lambda: [0, 0])
total = 0
for e in eng.train_prog:
    if e[0] is None:
        continue
    seen, nbytes = set(), 0
    for a in e[1]:
        if isinstance(a, int) and a in bufs and a not in seen:
            seen.add(a)
            nbytes += bufs[a][1]
    per_kernel[e[0].__name__][0] += 1
    per_kernel[e[0].__name__][1] += nbytes
    total += nbytes
peaks = {"hbm_gbs": 6454.6, "bf16_tflops": 1684.1}
try:
    peaks.update(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))))
except Exception:
    pass
flops = 3.0 * S.flops_forward(H, H) * B
out = {"batch": B, "size": H, "launches": sum(v[0] for v in per_kernel.values()), "activation_bytes_per_step": total,
       "hbm_roofline_ms": total / peaks["hbm_gbs"] / 1e6, "flops_per_step": flops,
       "tensor_roofline_ms": flops / peaks["bf16_tflops"] / 1e9,
       "per_kernel_mb": {k: [v[0], round(v[1] / 1e6, 1)] for k, v in sorted(per_kernel.items(), key=lambda kv: -kv[1][1])}}
print(json.dumps(out, indent=1))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/step_bytes.json", "w"), indent=1)
