#!/usr/bin/env python
"""Fused aggregation: correctness across ranks + bandwidth sweep (BASELINE.json config 5).

Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/agg_sweep.py [--sizes-mb 1,8,64,...]
Every rank publishes K client updates with known values; after the fused kernel every rank's global arena must equal
the NumPy weighted mean (fl_server.py:92-102 semantics), and the WHOLE arena is compared bit for bit across ranks.
Timing and roofline: crackfl/parallel/sweep.py (ranks aligned on the device by a barrier-only launch, CUDA events, max
over ranks; bytes per direction (partial_bytes + 4n)(G-1)/G over 900 GB/s, and over the measured 770 GB/s for context;
for G=1 the bound is HBM traffic over the measured copy bandwidth).
Works on CPU/Gloo too (host fallback path) so the logic is testable without GPUs.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from crackfl.parallel.runtime import OnBoxFederation, init_distributed
from crackfl.parallel.sweep import roofline_ms, time_aggregate


def same_on_all_ranks(t: torch.Tensor, chunk: int = 16 << 20) -> bool:
    """Bit-for-bit comparison of the whole tensor against rank 0's, in chunks (int32 view: NaN-safe, -0 != +0)."""
    ti = t.view(torch.int32)
    ok = True
    for o in range(0, ti.numel(), chunk):
        mine = ti[o:o + chunk]
        ref = mine.clone()
        dist.broadcast(ref, 0)
        ok = ok and bool(torch.equal(mine, ref))
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", default="1,8.233,64,256")
    ap.add_argument("--clients-per-gpu", type=int, default=1)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="gpurun_out/agg_sweep.json")
    ap.add_argument("--payloads", default="fp32,bf16_delta")
    a = ap.parse_args()
    init_distributed()
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    cuda = torch.cuda.is_available()
    results, ok_all = [], True
    K = a.clients_per_gpu
    for payload in a.payloads.split(","):
        for strategy in ("fedavg_weighted", "fedadam"):
            for mb in [float(x) for x in a.sizes_mb.split(",")]:
                if strategy == "fedadam" and mb > 64:
                    continue
                n = int(mb * 1e6 / 4) // 1024 * 1024
                fed = OnBoxFederation(clients_per_gpu=K, strategy=strategy, payload=payload, numel=n)
                dev = fed.global_arena.device
                idx = torch.arange(n, device=dev, dtype=torch.float32)
                glob = torch.sin(idx * 0.37) * 0.5                      # cheap deterministic data, built on the device
                fed.set_global(glob)
                ns_all = [[1 + ((r * K + k) % 5) for k in range(K)] for r in range(world)]

                def update_of(cid):
                    return glob + 0.01 * torch.cos(idx * (0.11 + 0.013 * cid))
                for k in range(K):
                    fed.publish(k, update_of(rank * K + k))
                fed.aggregate(ns_all[rank], weighted=True)
                # oracle (same arithmetic as fl_server.py:92-102 with sample-count weights), evaluated on `dev`
                flat_ns = np.array([x for row in ns_all for x in row], np.float64)
                coef = (flat_ns / flat_ns.sum()).astype(np.float32)
                avg = torch.zeros(n, device=dev)
                for cid, c in enumerate(coef):
                    avg += float(c) * update_of(cid)
                if strategy == "fedadam":
                    d = avg - glob
                    m = 0.1 * d; v = 0.01 * d * d
                    want = glob + 1e-2 * m / (v.sqrt() + 1e-3)
                else:
                    want = avg
                got = fed.global_arena.detach()
                tol = 3e-4 if payload == "bf16_delta" else 2e-6
                err = float((got - want).abs().max())
                ok = err < tol * (10 if strategy == "fedadam" else 1) + 1e-6
                del avg, want
                # all ranks bit-identical
                same = same_on_all_ranks(got) if world > 1 else True
                flag = torch.tensor([1.0 if (ok and same) else 0.0], device=fed.global_arena.device)
                if world > 1:
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                ok_all = ok_all and bool(flag.item() == 1.0)
                # timing
                if cuda:
                    tm = time_aggregate(fed, a.iters)
                else:
                    ts = sorted(fed.aggregate(ns_all[rank], weighted=True) for _ in range(a.iters))
                    tm = {"ms_best": ts[0], "ms_median": ts[len(ts) // 2]}
                roof = roofline_ms(n, world, payload, K)
                rec = {"payload": payload, "strategy": strategy, "mb": mb, "numel": n, "world": world, "clients_per_gpu": K,
                       "nvls": bool(getattr(fed, "nvls", False)), **tm, **roof,
                       "roofline_frac_900": roof["roofline_ms_900"] / tm["ms_best"] if tm["ms_best"] > 0 else None,
                       "roofline_frac_770": roof["roofline_ms_770"] / tm["ms_best"] if tm["ms_best"] > 0 else None,
                       "max_abs_err": err, "bit_identical_all_ranks_whole_arena": same, "ok": bool(flag.item() == 1.0)}
                results.append(rec)
                if rank == 0:
                    print(json.dumps(rec), flush=True)
                del fed
                if cuda:
                    torch.cuda.empty_cache()
    if rank == 0:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        with open(a.out, "w") as f:
            json.dump({"world": world, "cuda": cuda, "ok": ok_all, "results": results}, f, indent=1)
        print("AGG_SWEEP", "OK" if ok_all else "FAILED")
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0 if ok_all else 1


if __name__ == "__main__":
    sys.exit(main())
