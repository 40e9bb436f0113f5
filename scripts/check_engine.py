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
