"""Engine vs fp32 oracle at the other BASELINE.json resolutions / widths (loss, worst forward error, worst gradient
cosine), so that kernel changes tuned on 256x256 do not silently break the 128 / 448 / wide / odd configurations."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.models.spec import REFERENCE_SPEC, wide_spec
from crackfl.utils.check import engine_vs_oracle

ok = True
for name, spec, B, S in (("128", REFERENCE_SPEC, 16, 128), ("256", REFERENCE_SPEC, 4, 256), ("448", REFERENCE_SPEC, 2, 448),
                         ("wide2-128", wide_spec(2), 4, 128), ("odd-80x112", REFERENCE_SPEC, 3, (80, 112))):
    H, W = (S, S) if isinstance(S, int) else S
    rep = engine_vs_oracle(spec, B=B, H=H, W=W, conv_impl="tc", randomize_bn=False)
    wf = max(rep["fwd"].values())
    gc = {k: v[1] for k, v in rep["grad"].items() if not isinstance(v[0], str)}
    tail = {k: v for k, v in gc.items() if k.startswith(("conv2d_transpose_6", "conv2d_transpose_7", "conv2d_6", "conv2d_7"))}
    wg = min(gc.items(), key=lambda kv: kv[1])
    wt = min(tail.items(), key=lambda kv: kv[1]) if tail else ("-", 1.0)
    el, ol = rep["loss"]
    print(f"{name:12s} loss {el:.5f} vs {ol:.5f}   worst fwd rel err {wf:.4f}   worst grad cos {wg[1]:.4f} ({wg[0]})   decoder tail {wt[1]:.4f}", flush=True)
    ok &= abs(el - ol) < 0.02 * abs(ol) + 1e-3 and wf < 0.08 and wt[1] > 0.98
print("OK" if ok else "MISMATCH")
