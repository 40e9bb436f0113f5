#!/usr/bin/env python
"""3x3 weight gradient per decoder layer shape (256x256 step, B=16): per-(tap,chunk) kernel vs halo kernel.  CUDA events."""
import os, sys, json, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.ops import lib as L
lib = L.get_lib(); st = torch.cuda.current_stream().cuda_stream
B = 16
def timeit(fn, iters=8):
    best = 1e9
    for i in range(iters + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i >= 2: best = min(best, e0.elapsed_time(e1))
    return best * 1e3
for r, cin, cout in ((16, 256, 256), (32, 128, 128), (32, 128, 256), (64, 64, 64), (64, 64, 128), (128, 32, 32), (128, 32, 64)):
    x = torch.randn(B, r, r, cin, device="cuda").bfloat16(); dy = torch.randn(B, r, r, cout, device="cuda").bfloat16()
    dw = torch.zeros(9, cin, cout, device="cuda")
    t = timeit(lambda: L.check(lib.crackfl_wgrad_tc(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), B, r, r, cin, cout, 9, cin * cout, cout, 1, st)))
    print(os.environ.get("CRACKFL_WGRAD_HALO", "1"), r, cin, cout, round(t, 1), flush=True)
