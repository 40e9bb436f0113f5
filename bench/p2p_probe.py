#!/usr/bin/env python
"""Peer-memory bandwidth probe with the aggregation kernel itself (2 ranks): rank 0 pulls from / pushes to rank 1's
IPC-mapped arena while rank 1 idles, so read and write link bandwidth are measured separately and together."""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from crackfl.parallel.runtime import OnBoxFederation, init_distributed
from crackfl.ops import lib as L

init_distributed()
rank, world = dist.get_rank(), dist.get_world_size()
n = int(os.environ.get("PROBE_MB", "1024")) * 250000
fed = OnBoxFederation(clients_per_gpu=1, payload="fp32", numel=n)
lib = fed.lib
st = torch.cuda.current_stream()
local_src = torch.randn(n, device=fed.dev)
local_dst = torch.empty(n, device=fed.dev)
dist.barrier()
res = {}
if rank == 0:
    peer = 1
    cases = {
        "local_read+local_write": ([local_src.data_ptr()], [local_dst.data_ptr()]),
        "remote_read+local_write": ([fed.peer["partial"][peer]], [local_dst.data_ptr()]),
        "local_read+remote_write": ([local_src.data_ptr()], [fed.peer["global"][peer]]),
        "remote_read+remote_write": ([fed.peer["partial"][peer]], [fed.peer["global"][peer]]),
        "local+remote_read,local+remote_write": ([local_src.data_ptr(), fed.peer["partial"][peer]],
                                                   [local_dst.data_ptr(), fed.peer["global"][peer]]),
    }
    for unroll in (4, 8):
        for grid in (592, 1184, 4736):
            for name, (srcs, dsts) in cases.items():
                p = L.FedAggParams()
                for i, s in enumerate(srcs):
                    p.client[i] = s; p.coef[i] = 1.0 / len(srcs)
                for i, d in enumerate(dsts):
                    p.out[i] = d
                p.num_clients, p.num_out = len(srcs), len(dsts)
                p.begin, p.end = 0, n
                p.rank, p.world, p.epoch = 0, 1, 1
                p.unroll = unroll
                p.load_mode = int(os.environ.get("CRACKFL_AGG_LOAD_MODE", "0"))
                best = 1e9
                for it in range(4):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(st)
                    L.check(lib.crackfl_fedagg(C.byref(p), grid, st.cuda_stream))
                    e1.record(st); st.synchronize()
                    best = min(best, e0.elapsed_time(e1))
                gb = n * 4 / 1e9
                res[f"u{unroll}_g{grid}_{name}"] = {"ms": best, "GBps_per_stream": gb / best * 1e3}
                print(f"u{unroll} g{grid:5d} {name:40s} {best:8.3f} ms  {gb / best * 1e3:7.1f} GB/s per stream", flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/p2p_probe.json", "w"), indent=1)
dist.barrier()
dist.destroy_process_group()
