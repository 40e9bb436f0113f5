#!/usr/bin/env python
"""Fused aggregation: correctness across ranks + bandwidth sweep (BASELINE.json config 5).

Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/agg_sweep.py [--sizes-mb 1,8,64,...]
Every rank publishes K client updates with known values; after the fused kernel every rank's global arena must equal
the NumPy weighted mean (fl_server.py:92-102 semantics), bit-identical on all ranks.  Timing: CUDA events around the
launch(es), ranks released together by a barrier, max over ranks.  Roofline per GPU and direction: the G-1 shards of
its partial that the other owners pull PLUS its updated shard pushed to G-1 peers = (partial_bytes + 4n)(G-1)/G,
over the measured 770 GB/s per direction (B200_PROFILING.md; bench/p2p_probe.py measures 750-780 GB/s read and
~710 GB/s write with this very kernel); for G=1 the bound is HBM traffic over the measured copy bandwidth.
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
                ms = fed.aggregate(ns_all[rank], weighted=True)
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
                if world > 1:
                    chk = got[:4096].clone()
                    ref = chk.clone()
                    dist.broadcast(ref, 0)
                    same = bool(torch.equal(chk, ref))
                else:
                    same = True
                flag = torch.tensor([1.0 if (ok and same) else 0.0], device=fed.global_arena.device)
                if world > 1:
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                ok_all = ok_all and bool(flag.item() == 1.0)
                # timing
                times = []
                for _ in range(a.iters):
                    if cuda:
                        torch.cuda.synchronize()
                    if world > 1:
                        dist.barrier()   # start every rank together: the max-over-ranks time is then the kernel, not rank skew
                    times.append(fed.aggregate(ns_all[rank], weighted=True))
                times.sort()
                best, med = times[0], times[len(times) // 2]
                pbytes = n * (2 if payload == "bf16_delta" else 4)
                if world > 1:
                    # bytes that must cross NVLink per GPU per direction for pull-reduce + push-broadcast:
                    # outbound = my partial's shards served to the G-1 other owners + my updated shard pushed to G-1 peers
                    # (inbound is the mirror image); NVLink is full duplex, so the bound is one direction's bytes.
                    link_bytes = (pbytes + n * 4) * (world - 1) / world
                    roof_ms = link_bytes / 770e9 * 1e3
                else:
                    hbm = K * n * 4 + pbytes + pbytes + n * 4 * 2
                    roof_ms = hbm / 6454.6e9 * 1e3
                rec = {"payload": payload, "strategy": strategy, "mb": mb, "numel": n, "world": world, "clients_per_gpu": K,
                       "ms_best": best, "ms_median": med, "roofline_ms": roof_ms, "roofline_frac": roof_ms / best if best > 0 else None,
                       "max_abs_err": err, "ok": bool(flag.item() == 1.0)}
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
