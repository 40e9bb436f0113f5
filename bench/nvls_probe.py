#!/usr/bin/env python
"""Can torch's symmetric memory give us an NVLS multicast mapping on this box?  (2+ ranks)"""
import os, sys, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from crackfl.parallel.runtime import init_distributed
init_distributed()
rank = dist.get_rank()
try:
    import torch.distributed._symmetric_memory as symm
    n = 1 << 20
    t = symm.empty(n, dtype=torch.float32, device=torch.device("cuda", torch.cuda.current_device()))
    hdl = symm.rendezvous(t, group=dist.group.WORLD)
    info = {"rank": rank, "world": hdl.world_size, "buffer_ptrs": [hex(p) for p in hdl.buffer_ptrs],
            "signal_pad_ptrs": [hex(p) for p in hdl.signal_pad_ptrs], "multicast_ptr": hex(hdl.multicast_ptr),
            "has_multicast": bool(getattr(hdl, "multicast_ptr", 0))}
    print("NVLS_PROBE", info, flush=True)
except Exception:
    print("NVLS_PROBE_FAILED rank", rank, traceback.format_exc()[-1500:], flush=True)
dist.barrier()
dist.destroy_process_group()
