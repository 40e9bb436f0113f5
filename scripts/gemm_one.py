#!/usr/bin/env python
"""One GEMM-shaped launch for ncu: python scripts/gemm_one.py kind H Cin N taps   (kind: fwd | wgrad | dwbwd | bnapply)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from crackfl.ops import lib as L
lib = L.get_lib(); st = torch.cuda.current_stream().cuda_stream
kind, r, cin, n, taps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
B = 16
x = torch.randn(B, r, r, cin, device="cuda").bfloat16()
if kind == "fwd":
    wk = (torch.randn(n, taps * cin, device="cuda") / np.sqrt(taps * cin)).bfloat16()
    bias = torch.randn(n, device="cuda"); y = torch.zeros(B, r, r, n, device="cuda", dtype=torch.bfloat16); stats = torch.zeros(2 * n, device="cuda")
    fn = lambda: L.check(lib.crackfl_conv_tc_fwd(x.data_ptr(), wk.data_ptr(), bias.data_ptr(), 0, y.data_ptr(), stats.data_ptr(), B, r, r, cin, n, taps, st))
elif kind == "wgrad":
    dy = torch.randn(B, r, r, n, device="cuda").bfloat16(); dw = torch.zeros(taps, cin, n, device="cuda")
    fn = lambda: L.check(lib.crackfl_wgrad_tc_ws_cv(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), B, r, r, cin, n, taps, cin * n, n, 1, 0, 0, 0, st))
elif kind == "dwbwd":
    g = torch.randn(B, r, r, cin, device="cuda").bfloat16(); w = torch.randn(9, cin, device="cuda")
    bn = torch.cat([torch.rand(cin, device="cuda") + 0.5, torch.randn(cin, device="cuda") * 0.1, torch.zeros(cin, device="cuda"), torch.ones(cin, device="cuda")]).contiguous()
    dz = torch.empty_like(x); dwk = torch.zeros(9, cin, device="cuda"); sums = torch.zeros(2 * cin, device="cuda")
    fn = lambda: L.check(lib.crackfl_dw3x3_bwd_tma_bn(g.data_ptr(), x.data_ptr(), w.data_ptr(), bn.data_ptr(), 0, dz.data_ptr(), dwk.data_ptr(), B, r, r, cin, 1, 0, bn.data_ptr(), sums.data_ptr(), st))
elif kind == "bnapply":
    g = torch.randn(B, r, r, cin, device="cuda").bfloat16()
    bn = torch.cat([torch.rand(cin, device="cuda") + 0.5, torch.randn(cin, device="cuda") * 0.1, torch.zeros(cin, device="cuda"), torch.ones(cin, device="cuda")]).contiguous()
    sums = torch.zeros(2 * cin, device="cuda"); out = torch.empty_like(x); dg = torch.zeros(cin, device="cuda"); db = torch.zeros(cin, device="cuda")
    fn = lambda: L.check(lib.crackfl_bn_bwd_apply(g.data_ptr(), 0, x.data_ptr(), bn.data_ptr(), sums.data_ptr(), out.data_ptr(), dg.data_ptr(), db.data_ptr(), B, r, r, cin, 0, st))
elif kind == "dwfwd":
    w = torch.randn(9, cin, device="cuda"); out = torch.empty_like(x)
    bn = torch.cat([torch.rand(cin, device="cuda") + 0.5, torch.randn(cin, device="cuda") * 0.1]).contiguous()
    fn = lambda: L.check(lib.crackfl_dw3x3_tma(x.data_ptr(), w.data_ptr(), bn.data_ptr(), None, out.data_ptr(), 0, 0, 0, B, r, r, cin, 1, 0, 0, 0, st))
elif kind == "pooladd":
    bn = torch.cat([torch.rand(cin, device="cuda") + 0.5, torch.randn(cin, device="cuda") * 0.1]).contiguous()
    res = torch.randn(B, r // 2, r // 2, cin, device="cuda").bfloat16(); out = torch.empty_like(res); even = torch.empty(B, r // 4, r // 4, cin, device="cuda", dtype=torch.bfloat16)
    arg = torch.empty(B, r // 2, r // 2, cin, device="cuda", dtype=torch.uint8)
    fn = lambda: L.check(lib.crackfl_pool_add_fwd(x.data_ptr(), bn.data_ptr(), res.data_ptr(), out.data_ptr(), even.data_ptr(), 0, arg.data_ptr(), B, r, r, cin, st))
elif kind == "poolbwd":
    bn = torch.cat([torch.rand(cin, device="cuda") + 0.5, torch.randn(cin, device="cuda") * 0.1, torch.zeros(cin, device="cuda"), torch.ones(cin, device="cuda")]).contiguous()
    g = torch.randn(B, r // 2, r // 2, cin, device="cuda").bfloat16(); arg = torch.randint(0, 9, (B, r // 2, r // 2, cin), device="cuda", dtype=torch.uint8)
    dz = torch.empty_like(x); sums = torch.zeros(2 * cin, device="cuda")
    fn = lambda: L.check(lib.crackfl_pool_bwd_bn(g.data_ptr(), arg.data_ptr(), x.data_ptr(), bn.data_ptr(), dz.data_ptr(), sums.data_ptr(), B, r, r, cin, st))
elif kind == "down2sum":
    out = torch.empty(B, r // 2, r // 2, cin, device="cuda", dtype=torch.bfloat16)
    fn = lambda: L.check(lib.crackfl_down2sum(x.data_ptr(), 0, 0, out.data_ptr(), B, r // 2, r // 2, cin, st))
elif kind == "colsum":
    out = torch.zeros(cin, device="cuda")
    fn = lambda: L.check(lib.crackfl_colsum(x.data_ptr(), out.data_ptr(), B * r * r, cin, st))
elif kind == "bnreduce":
    g = torch.randn(B, r, r, cin, device="cuda").bfloat16()
    bn = torch.cat([torch.rand(cin, device="cuda") + 0.5, torch.randn(cin, device="cuda") * 0.1, torch.zeros(cin, device="cuda"), torch.ones(cin, device="cuda")]).contiguous()
    sums = torch.zeros(2 * cin, device="cuda")
    fn = lambda: L.check(lib.crackfl_bn_bwd_reduce(g.data_ptr(), 0, x.data_ptr(), bn.data_ptr(), sums.data_ptr(), B, r, r, cin, 0, st))
for _ in range(3):
    fn()
torch.cuda.synchronize()
print("ok")
