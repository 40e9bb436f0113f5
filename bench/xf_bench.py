#!/usr/bin/env python
"""Per-layer timing of the transform-prologue halo kernel (conv_tc_xf_k) against the kernels it replaces, at the shapes of
the 256x256, B=16 step: depthwise forward (TMA stencil vs tensor-core block-diagonal) and the decoder's second conv
(bn_act pass + halo conv vs fused prologue).  CUDA events, best of N, inputs warm in L2 as in the step."""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from crackfl.ops import lib as L
lib = L.get_lib()
st = torch.cuda.current_stream().cuda_stream
B = 16


def timeit(fn, iters=8):
    best = 1e9
    for i in range(iters + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i >= 2:
            best = min(best, e0.elapsed_time(e1))
    return best * 1e3


rows = []
for r, c in ((128, 32), (128, 64), (64, 64), (64, 128), (32, 128), (32, 256)):
    x = torch.randn(B, r, r, c, device="cuda").bfloat16(); out = torch.empty_like(x)
    w = torch.randn(9, c, device="cuda")
    bn = torch.cat([torch.rand(c, device="cuda") + 0.5, torch.randn(c, device="cuda") * 0.1]).contiguous()
    kc = 64 if c % 64 == 0 else 32
    ops = torch.zeros((c // kc) * 9 * kc * kc, device="cuda", dtype=torch.bfloat16)
    d = L.DwDiagDesc(src_off=0, dst_off=0, C=c, kc=kc, flip=0, pad_=0)
    desc = torch.frombuffer(bytearray(bytes(d)), dtype=torch.uint8).clone().cuda()
    L.check(lib.crackfl_dw_diag_prep(w.data_ptr(), ops.data_ptr(), desc.data_ptr(), 1, st))
    t_old = timeit(lambda: L.check(lib.crackfl_dw3x3_tma(x.data_ptr(), w.data_ptr(), bn.data_ptr(), None, out.data_ptr(), 0, 0, 0, B, r, r, c, 1, 0, 0, 0, st)))
    t_new = timeit(lambda: L.check(lib.crackfl_conv_tc_xf(x.data_ptr(), ops.data_ptr(), 0, 0, out.data_ptr(), 0, B, r, r, c, c, 2, bn.data_ptr(), None, 0, 1, st)))
    t_relu = timeit(lambda: L.check(lib.crackfl_conv_tc_xf(x.data_ptr(), ops.data_ptr(), 0, 0, out.data_ptr(), 0, B, r, r, c, c, 1, bn.data_ptr(), None, 0, 1, st)))
    t_none = timeit(lambda: L.check(lib.crackfl_conv_tc_xf(x.data_ptr(), ops.data_ptr(), 0, 0, out.data_ptr(), 0, B, r, r, c, c, 3, bn.data_ptr(), None, 0, 1, st)))
    mb = 2 * x.numel() * 2 / 1e6
    rows.append({"op": "dw_fwd_bn_relu", "H": r, "C": c, "stencil_us": t_old, "tensor_core_us": t_new, "tc_relu_only_us": t_relu, "tc_no_transform_us": t_none, "hbm_floor_us": mb / 6550.7 * 1e3})
    print(rows[-1], flush=True)
for r, c in ((16, 256), (32, 128), (64, 64), (128, 32)):
    x = torch.randn(B, r, r, c, device="cuda").bfloat16(); a1 = torch.empty_like(x); y = torch.empty_like(x)
    wk = (torch.randn(c, 9 * c, device="cuda") / np.sqrt(9 * c)).bfloat16()
    bias = torch.randn(c, device="cuda"); stats = torch.zeros(2 * c, device="cuda")
    bn = torch.cat([torch.rand(c, device="cuda") + 0.5, torch.randn(c, device="cuda") * 0.1]).contiguous()

    def old():
        L.check(lib.crackfl_bn_act(x.data_ptr(), bn.data_ptr(), a1.data_ptr(), 0, B, r, r, c, 1, st))
        L.check(lib.crackfl_conv_tc_fwd(a1.data_ptr(), wk.data_ptr(), bias.data_ptr(), 0, y.data_ptr(), stats.data_ptr(), B, r, r, c, c, 9, st))
    t_old = timeit(old)
    t_new = timeit(lambda: L.check(lib.crackfl_conv_tc_xf(x.data_ptr(), wk.data_ptr(), bias.data_ptr(), 0, y.data_ptr(), stats.data_ptr(), B, r, r, c, c, 2, bn.data_ptr(), None, a1.data_ptr(), 0, st)))
    rows.append({"op": "dec_ct2", "H": r, "C": c, "bn_act+conv_us": t_old, "fused_us": t_new})
    print(rows[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/xf_bench.json", "w"), indent=1)
