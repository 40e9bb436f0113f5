#!/usr/bin/env python
"""Aggregation-kernel timing points (aligned ranks, see crackfl/parallel/sweep.py).
torchrun --nproc-per-node N bench/agg_points.py --sizes-mb 1,8.233,64,256,1024,4096 [--out file.json]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from crackfl.parallel.runtime import init_distributed
from crackfl.parallel.sweep import sweep_points

ap = argparse.ArgumentParser()
ap.add_argument("--sizes-mb", default="1,8.233,64,256,1024")
ap.add_argument("--payload", default="fp32")
ap.add_argument("--strategy", default="fedavg")
ap.add_argument("--iters", type=int, default=8)
ap.add_argument("--out", default="")
ap.add_argument("--tag", default="")
a = ap.parse_args()
init_distributed()
rank = dist.get_rank() if dist.is_initialized() else 0
pts = sweep_points([float(x) for x in a.sizes_mb.split(",")], payload=a.payload, strategy=a.strategy, iters=a.iters)
if rank == 0:
    for p in pts:
        print(json.dumps({"tag": a.tag, **{k: (round(v, 5) if isinstance(v, float) else v) for k, v in p.items()}}), flush=True)
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump({"tag": a.tag, "env": {k: v for k, v in os.environ.items() if k.startswith("CRACKFL_")}, "points": pts},
                  open(a.out, "w"), indent=1)
if dist.is_initialized():
    dist.barrier(); dist.destroy_process_group()
