#!/usr/bin/env python
"""Algorithmic-minimum HBM traffic of one train step (CPU only, from the model spec) - the number the schedule's own
traffic (scripts/step_bytes.py) is compared against in profiles/ROOFLINE.md.

Model (stated so it can be argued with).  Training-mode BatchNorm puts a grid-wide dependency behind every BN'd conv
output z (its statistics need the whole tensor) and, in the backward pass, behind every dz (BN backward needs
sum(dz) and sum(dz*xhat) before any dy can be formed).  Everything else - ReLU, BN-apply, nearest upsample, residual add,
max-pool, the depthwise->pointwise pair of a separable conv, the loss head - is elementwise or tile-local and is assumed
to be fused into a producer's epilogue or a consumer's load, with elementwise values recomputed instead of stored.
Convolutions are not recomputed.  Low-resolution decoder tensors (nearest upsample commutes with 1x1 conv / BN / add) are
counted at low resolution.  bf16 activations, weights/gradients/statistics (a few MB) ignored.  Then:
  forward : every z written once and read once (by the consumer that applies BN on load); every block output (the
            pre-activation sum that feeds the next block AND its residual branch) written once, read twice
  backward: every z read twice more (BN-backward reduce phase, then apply phase fused into dgrad+wgrad); every dz
            written once and read twice (same two phases); every block-output gradient written once and read once;
            the conv inputs that the weight gradients need (act(bn(z)) / block outputs, recomputed on load) read once more
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from crackfl.models.spec import REFERENCE_SPEC as S

B, H = int(os.environ.get("B", 16)), int(os.environ.get("SIZE", 256))
hbm = 6454.6
try:
    hbm = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
bf = 2
z, blocks = [], []
r = H // 2
z.append(("entry", B * r * r * S.entry * bf))
for f in S.down:
    z += [(f"down{f}.z1", B * r * r * f * bf), (f"down{f}.z2", B * r * r * f * bf)]
    r //= 2
    blocks.append((f"down{f}.out", B * r * r * f * bf))
for f in S.up:
    z += [(f"up{f}.z1", B * r * r * f * bf), (f"up{f}.z2", B * r * r * f * bf)]
    blocks.append((f"up{f}.out(lowres)", B * r * r * f * bf))
    r *= 2
Z, X = sum(b for _, b in z), sum(b for _, b in blocks)
inp = B * H * H * 3 + B * H * H  # uint8 image + mask
fwd = inp + 2 * Z + 3 * X
bwd = 2 * Z + 3 * Z + 2 * X + (Z + X)
tot = fwd + bwd
out = {"batch": B, "size": H, "z_bytes": Z, "block_output_bytes": X, "forward_bytes": fwd, "backward_bytes": bwd,
       "min_bytes_per_step": tot, "hbm_gbs_measured": hbm, "min_hbm_ms": tot / hbm / 1e6}
print(json.dumps(out, indent=1))
