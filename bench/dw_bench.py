#!/usr/bin/env python
"""Timing of the memory-bound kernels (depthwise 3x3 fwd / dgrad / wgrad, BN backward, pooling tails) at the shapes of
the 256x256, B=16 U-Net step.  CUDA events, best of N; by default the inputs stay warm in L2 as they are in the step
(producer -> consumer adjacency), ``--flush`` writes 256 MB between launches.  Reports achieved bytes/s against the
measured HBM copy bandwidth (MEASURED_PEAKS.json)."""
import argparse, ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.ops import lib as L

ap = argparse.ArgumentParser()
ap.add_argument("--only", default="")
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--size", type=int, default=256)
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--flush", action="store_true")
ap.add_argument("--shape", default="", help="H,C: only this shape")
ap.add_argument("--out", default="gpurun_out/dw_bench.json")
a = ap.parse_args()
lib = L.get_lib()
B, S = a.batch, a.size
peaks = {"hbm_gbs": 6454.6}
try:
    peaks.update(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))))
except Exception:
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream


def timeit(fn):
    best = 1e9
    for i in range(a.iters + 1):
        if a.flush:
            flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i > 0:
            best = min(best, e0.elapsed_time(e1))
    return best


rows = []
for r, c in ((S // 2, 32), (S // 2, 64), (S // 4, 64), (S // 4, 128), (S // 8, 128), (S // 8, 256)):
    if a.shape and a.shape != f"{r},{c}":
        continue
    x = torch.randn(B, r, r, c, device="cuda").bfloat16()
    g = torch.randn(B, r, r, c, device="cuda").bfloat16()
    out = torch.empty_like(x)
    w = torch.randn(9, c, device="cuda")
    bn = torch.cat([torch.rand(c, device="cuda") + 0.5, torch.randn(c, device="cuda") * 0.1])
    dwk = torch.zeros(9, c, device="cuda")
    sums = torch.zeros(2 * c, device="cuda")
    bnfull = torch.cat([bn, torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")])
    dg, db = torch.zeros(c, device="cuda"), torch.zeros(c, device="cuda")
    n = x.numel() * 2
    cases = (
        ("dw_fwd_relu", lambda: lib.crackfl_dw3x3_tma(x.data_ptr(), w.data_ptr(), 0, None, out.data_ptr(), 0, 0, 0, B, r, r, c, 0, 0, 0, 0, st), 2 * n),
        ("dw_fwd_bn", lambda: lib.crackfl_dw3x3_tma(x.data_ptr(), w.data_ptr(), bn.data_ptr(), None, out.data_ptr(), 0, 0, 0, B, r, r, c, 1, 0, 0, 0, st), 2 * n),
        ("dw_dgrad_bn", lambda: lib.crackfl_dw3x3_tma(g.data_ptr(), w.data_ptr(), 0, None, out.data_ptr(), x.data_ptr(), bn.data_ptr(), 0, B, r, r, c, 2, 1, 1, 0, st), 3 * n),
        ("dw_wgrad_bn", lambda: lib.crackfl_dw3x3_wgrad_tma(g.data_ptr(), x.data_ptr(), bn.data_ptr(), dwk.data_ptr(), B, r, r, c, 1, st), 2 * n),
        ("dw_bwd_merged", lambda: lib.crackfl_dw3x3_bwd_tma(g.data_ptr(), x.data_ptr(), w.data_ptr(), bn.data_ptr(), 0, out.data_ptr(), dwk.data_ptr(), B, r, r, c, 1, 0, st), 3 * n),
        ("bn_bwd_reduce", lambda: lib.crackfl_bn_bwd_reduce(g.data_ptr(), 0, x.data_ptr(), bnfull.data_ptr(), sums.data_ptr(), B, r, r, c, 0, st), 2 * n),
        ("bn_bwd_apply", lambda: lib.crackfl_bn_bwd_apply(g.data_ptr(), 0, x.data_ptr(), bnfull.data_ptr(), sums.data_ptr(), out.data_ptr(), dg.data_ptr(), db.data_ptr(), B, r, r, c, 0, st), 3 * n),
    )
    for name, fn, byt in cases:
        if a.only and a.only not in name:
            continue
        ms = timeit(lambda: L.check(fn()))
        rec = {"kernel": name, "H": r, "C": c, "us": ms * 1e3, "min_mb": byt / 1e6, "gbs": byt / ms / 1e6,
               "frac_hbm_measured": byt / ms / 1e6 / peaks["hbm_gbs"]}
        rows.append(rec)
        print(f"{name:14s} H={r:4d} C={c:4d} {ms*1e3:8.1f} us  {rec['gbs']:8.0f} GB/s  {rec['frac_hbm_measured']:.2f} of HBM copy", flush=True)
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
json.dump({"peaks": peaks, "flush": a.flush, "rows": rows}, open(a.out, "w"), indent=1)
