#!/usr/bin/env python
"""Per-layer timing of the GEMM-shaped kernels (tcgen05 conv fwd/dgrad and wgrad) at the shapes of the 256x256, B=16
U-Net step.  CUDA events, L2 flushed between timed launches (a 256 MB write), best of N.  Reports achieved
bytes/s against the measured HBM copy bandwidth and FLOP/s against the measured bf16 GEMM peak (MEASURED_PEAKS.json)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.ops import lib as L

ap = argparse.ArgumentParser()
ap.add_argument("--only", default="")
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--size", type=int, default=256)
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--out", default="gpurun_out/conv_bench.json")
ap.add_argument("--no-flush", action="store_true")
a = ap.parse_args()
lib = L.get_lib()
B, S = a.batch, a.size
R = [S // 2, S // 4, S // 8, S // 16]
shapes = []  # (name, H, Cin, N, taps)
# encoder pointwise / residual
for k, (cin, f) in enumerate([(32, 64), (64, 128), (128, 256)]):
    shapes += [(f"e{k}.pw1", R[k], cin, f, 1), (f"e{k}.pw2", R[k], f, f, 1), (f"e{k}.res", R[k] // 2, cin, f, 1)]
for j, (cin, f) in enumerate([(256, 256), (256, 128), (128, 64), (64, 32)]):
    r = R[3 - j]
    shapes += [(f"d{j}.ct1", r, cin, f, 9), (f"d{j}.ct2", r, f, f, 9), (f"d{j}.res", r if j == 0 else r // 2, cin, f, 1)]
peaks = {"hbm_gbs": 6454.6, "bf16_tflops": 1684.1}
try:
    peaks.update(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))))
except Exception:
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream


def timeit(fn):
    best = 1e9
    for i in range(a.iters + 1):
        if not a.no_flush:
            flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i > 0:
            best = min(best, e0.elapsed_time(e1))
    return best


rows = []
for name, r, cin, n, taps in shapes:
    if a.only and a.only not in name:
        continue
    x = torch.randn(B, r, r, cin, device="cuda").bfloat16()
    dy = torch.randn(B, r, r, n, device="cuda").bfloat16()
    w = (torch.randn(n, taps * cin, device="cuda") / (taps * cin) ** 0.5).bfloat16()
    wd = (torch.randn(cin, taps * n, device="cuda") / (taps * n) ** 0.5).bfloat16()
    y = torch.empty(B, r, r, n, device="cuda", dtype=torch.bfloat16)
    dx = torch.empty(B, r, r, cin, device="cuda", dtype=torch.bfloat16)
    dw = torch.zeros(taps, cin, n, device="cuda")
    stats = torch.zeros(2 * n, device="cuda")
    M = B * r * r
    flops = 2.0 * M * cin * n * taps
    for kind, fn, byt in (
        ("fwd", lambda: L.check(lib.crackfl_conv_tc_fwd(x.data_ptr(), w.data_ptr(), 0, 0, y.data_ptr(), stats.data_ptr(), B, r, r, cin, n, taps, st)), (M * cin + M * n) * 2),
        ("fwd_nostat", lambda: L.check(lib.crackfl_conv_tc_fwd(x.data_ptr(), w.data_ptr(), 0, 0, y.data_ptr(), 0, B, r, r, cin, n, taps, st)), (M * cin + M * n) * 2),
        ("dgrad", lambda: L.check(lib.crackfl_conv_tc_fwd(dy.data_ptr(), wd.data_ptr(), 0, 0, dx.data_ptr(), 0, B, r, r, n, cin, taps, st)), (M * cin + M * n) * 2),
        ("wgrad", lambda: L.check(lib.crackfl_wgrad_tc(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), B, r, r, cin, n, taps, cin * n, n, 1, st)), (M * cin + M * n) * 2),
    ):
        ms = timeit(fn)
        rec = {"layer": name, "kind": kind, "H": r, "Cin": cin, "N": n, "taps": taps, "ms": ms, "gflop": flops / 1e9,
               "tflops": flops / ms / 1e9, "frac_bf16_peak_measured": flops / ms / 1e9 / peaks["bf16_tflops"],
               "min_mb": byt / 1e6, "gbs": byt / ms / 1e6, "frac_hbm_measured": byt / ms / 1e6 / peaks["hbm_gbs"]}
        rows.append(rec)
        print(f"{name:8s} {kind:6s} H={r:4d} Cin={cin:4d} N={n:4d} t={taps}  {ms*1e3:8.1f} us  {rec['tflops']:7.1f} TF/s ({100*rec['frac_bf16_peak_measured']:4.1f}%)  "
              f"{rec['gbs']:7.0f} GB/s ({100*rec['frac_hbm_measured']:4.1f}% of HBM copy)", flush=True)
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
json.dump({"peaks": peaks, "rows": rows}, open(a.out, "w"), indent=1)
tot = {}
for r_ in rows:
    tot[r_["kind"]] = tot.get(r_["kind"], 0) + r_["ms"]
print("totals (ms):", tot)
