#!/usr/bin/env python
"""Where the fixed cost of the aggregation launch goes: barrier-only launch vs tiny / small shards (aligned ranks)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from crackfl.parallel.runtime import OnBoxFederation, init_distributed
from crackfl.parallel.sweep import time_aggregate

init_distributed()
rank = dist.get_rank() if dist.is_initialized() else 0
out = {}
fed = OnBoxFederation(clients_per_gpu=1, numel=1 << 20)
fed.set_global(torch.zeros(1 << 20)); fed.configure([1.0], weighted=False)
st = torch.cuda.current_stream()
ts = []
for i in range(12):
    fed.align()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); fed.align(); e1.record(st); e1.synchronize()
    ts.append(e0.elapsed_time(e1))
t = torch.tensor(sorted(ts[2:]), device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
out["barrier_only_ms_best_med"] = [float(t[0]), float(t[len(t) // 2])]
del fed
for mb in (0.016, 0.262, 1.0, 4.0, 8.233):
    n = max(4096, int(mb * 1e6 / 4) // 1024 * 1024)
    fed = OnBoxFederation(clients_per_gpu=1, numel=n)
    fed.set_global(torch.zeros(n)); fed.configure([1.0], weighted=False)
    out[f"{mb}MB"] = time_aggregate(fed, 10)
    del fed
if rank == 0:
    print(json.dumps(out))
dist.barrier(); dist.destroy_process_group()
