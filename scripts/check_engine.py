"""GPU diagnostic: engine vs oracle error tables for every tap and gradient -> gpurun_out/."""
import json, os, sys, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crackfl.utils.check import engine_vs_oracle, format_report

os.makedirs("gpurun_out", exist_ok=True)
impls = sys.argv[1:] or ["simt"]
for impl in impls:
    for (B, H, W) in [(4, 64, 64), (2, 128, 128)]:
        try:
            rep = engine_vs_oracle(B=B, H=H, W=W, conv_impl=impl)
            txt = format_report(rep)
        except Exception:
            txt = traceback.format_exc()
        hdr = f"==== impl={impl} B={B} H={H} W={W}"
        print(hdr); print(txt)
        with open(f"gpurun_out/check_engine_{impl}.txt", "a") as f:
            f.write(hdr + "\n" + txt + "\n")

from crackfl.utils.check import directional_fd_check, bf16_oracle_noise
try:
    fd = directional_fd_check(conv_impl=impls[0])
    noise = bf16_oracle_noise()
    txt = "-- directional finite differences (engine loss along engine gradient)\n" + "\n".join(
        f"  {k:44s} |g| {v['engine_|g|']:.4e}  fd {v['fd_slope']:.4e}  oracle d.g {v['oracle_d.g']:.4e}" for k, v in fd.items())
    txt += "\n-- plain-PyTorch bf16 vs fp32 gradient cosine (noise floor of the comparison)\n" + "\n".join(
        f"  {k:44s} {v:.5f}" for k, v in noise.items())
except Exception:
    txt = traceback.format_exc()
print(txt)
with open(f"gpurun_out/check_engine_{impls[0]}.txt", "a") as f:
    f.write(txt + "\n")
