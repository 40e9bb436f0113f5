"""Timing of the fused aggregation kernel at a given model size (BASELINE.json config 5), shared by ``bench.py`` and
``bench/agg_sweep.py``.

How it is timed: every rank first enqueues ``fed.align()`` (a barrier-only launch), so that the aggregation launch behind
it starts within microseconds of its peers', then CUDA events around ``fed.aggregate()`` on the launching stream; the
figure is the MAX over ranks of each rank's best iteration.  Without the alignment the events would mostly measure how
far apart the ranks' hosts reached the launch.

Roofline: bytes that must cross NVLink per GPU and direction for pull-reduce + push-broadcast over peer memory =
(partial_bytes + 4n)(G-1)/G (the G-1 shards of this rank's partial served to their owners plus this rank's updated shard
stored to G-1 peers; inbound is the mirror image and the links are full duplex), over 900 GB/s (the NVLink 5 rate per
direction) and, for context, over the 770 GB/s a plain peer copy reaches on this pool (B200_PROFILING.md).
"""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.distributed as dist

LINK_GBS_NOMINAL = 900.0
LINK_GBS_MEASURED = 770.0
HBM_GBS_MEASURED = 6454.6


def time_aggregate(fed, iters: int = 8, warmup: int = 2) -> dict:
    """Best / median device time (ms) of ``fed.aggregate()``, max over ranks.  ``fed`` must be configured."""
    times = []
    st = torch.cuda.current_stream(fed.dev)
    for it in range(warmup + iters):
        fed.align()
        fed.aggregate()
        ms = fed.last_agg_ms
        if it >= warmup:
            times.append(ms)
    fed.check()
    t = torch.tensor(times, device=fed.dev)
    if fed.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # per-iteration max over ranks
    ts = sorted(float(x) for x in t.tolist())
    return {"ms_best": ts[0], "ms_median": ts[len(ts) // 2]}


def roofline_ms(n: int, world: int, payload: str, clients_per_gpu: int = 1) -> dict:
    pbytes = n * (2 if payload == "bf16_delta" else 4)
    if world > 1:
        link_bytes = (pbytes + n * 4) * (world - 1) / world
        return {"link_bytes_per_dir": link_bytes, "roofline_ms_900": link_bytes / (LINK_GBS_NOMINAL * 1e9) * 1e3,
                "roofline_ms_770": link_bytes / (LINK_GBS_MEASURED * 1e9) * 1e3}
    hbm = clients_per_gpu * n * 4 + pbytes + pbytes + n * 4 * 2
    r = hbm / (HBM_GBS_MEASURED * 1e9) * 1e3
    return {"hbm_bytes": hbm, "roofline_ms_900": r, "roofline_ms_770": r}


def sweep_points(sizes_mb: Sequence[float], payload: str = "fp32", strategy: str = "fedavg", iters: int = 8,
                 nvls=None) -> List[dict]:
    """One timed point per size: a fresh federation of ``numel = size/4`` synthetic parameters, one client per GPU."""
    from .runtime import OnBoxFederation
    out = []
    for mb in sizes_mb:
        n = int(mb * 1e6 / 4) // 1024 * 1024
        fed = OnBoxFederation(clients_per_gpu=1, strategy=strategy, payload=payload, numel=n, nvls=nvls)
        idx = torch.arange(n, device=fed.dev, dtype=torch.float32)
        fed.set_global(torch.sin(idx * 0.37) * 0.5)
        del idx
        fed.configure([1.0], weighted=False)
        rec = {"mb": mb, "numel": n, "world": fed.world, "payload": payload, "strategy": strategy, "nvls": bool(fed.nvls)}
        rec.update(time_aggregate(fed, iters))
        rec.update(roofline_ms(n, fed.world, payload))
        rec["roofline_frac_900"] = rec["roofline_ms_900"] / rec["ms_best"]
        rec["roofline_frac_770"] = rec["roofline_ms_770"] / rec["ms_best"]
        out.append(rec)
        del fed
        torch.cuda.empty_cache()
    return out
