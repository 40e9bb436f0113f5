"""Compare the tensor-core entry conv (im2col + tcgen05 GEMMs) against the CUDA-core entry kernels: forward y0 and the
kernel gradient of the same step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.models.unet import UNetEngine
from crackfl.data.generator import synth_crack_batch

B, H = 8, 64
x, y = synth_crack_batch(B, H, H, 0)
res = {}
for tc in ("0", "1"):
    os.environ["CRACKFL_ENTRY_TC"] = tc
    eng = UNetEngine(S, B, H, H, cuda_graph=False)
    eng.load_weights(S.init_weights(0)); eng.reset_optimizer(); eng.set_batch(x, y)
    eng.zero_region.zero_(); eng.grads.zero_(); eng._region_dirty = True
    eng._run(eng.grad_prog)
    torch.cuda.synchronize()
    t = S.tensors[S.index("conv2d/kernel")]
    res[tc] = (eng._bufs["y0"].float().clone(), eng.grads[t.offset:t.offset + t.numel].clone(),
               eng.grads.clone())
y0a, ga, fa = res["0"]; y0b, gb, fb = res["1"]
cos = lambda a, b: float((a * b).sum() / (a.norm() * b.norm() + 1e-30))
print("y0 rel err", float((y0a - y0b).norm() / y0a.norm()))
print("conv2d/kernel grad cos", cos(ga, gb), "norms", float(ga.norm()), float(gb.norm()))
print("all grads cos", cos(fa, fb))
