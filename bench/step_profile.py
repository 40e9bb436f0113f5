#!/usr/bin/env python
"""In-situ kernel timing of one train step (CUDA events between launches, eager, single stream)."""
import collections, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.data.synthetic_device import synth_crack_batch_device
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.models.unet import UNetEngine

B, H = int(os.environ.get("B", 16)), int(os.environ.get("SIZE", 256))
eng = UNetEngine(S, B, H, H, cuda_graph=False)
eng.load_weights(S.init_weights(0)); eng.reset_optimizer()
x, y = synth_crack_batch_device(B, H, H, 1)
eng.set_batch(x, y)
for _ in range(3):
    eng.train_step()
rows = eng.profile_step(5)
tot = sum(t for _, t in rows)
agg = collections.defaultdict(lambda: [0, 0.0])
for n, t in rows:
    agg[n][0] += 1; agg[n][1] += t
print(f"sum of per-launch event intervals: {tot:.3f} ms over {len(rows)} launches")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:28s} n={c:3d} {t*1e3:9.1f} us {100*t/tot:5.1f}%")
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"total_ms": tot, "launches": [{"i": i, "name": n, "us": t * 1e3} for i, (n, t) in enumerate(rows)]},
          open("gpurun_out/step_profile.json", "w"), indent=0)
