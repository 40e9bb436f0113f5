#!/usr/bin/env python
"""One shape of the tensor-core depthwise / fused-prologue kernel, for ncu: python scripts/xf_one.py H C diag"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from crackfl.ops import lib as L
lib = L.get_lib(); st = torch.cuda.current_stream().cuda_stream
r, c, diag = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
B = 16
x = torch.randn(B, r, r, c, device="cuda").bfloat16(); out = torch.empty_like(x)
bn = torch.cat([torch.rand(c, device="cuda") + 0.5, torch.randn(c, device="cuda") * 0.1]).contiguous()
if diag:
    kc = 64 if c % 64 == 0 else 32
    wk = torch.zeros((c // kc) * 9 * kc * kc, device="cuda", dtype=torch.bfloat16)
else:
    wk = (torch.randn(c, 9 * c, device="cuda") / np.sqrt(9 * c)).bfloat16()
for _ in range(3):
    L.check(lib.crackfl_conv_tc_xf(x.data_ptr(), wk.data_ptr(), 0, 0, out.data_ptr(), 0, B, r, r, c, c, 2, bn.data_ptr(), None, 0, diag, st))
torch.cuda.synchronize()
print("ok")
