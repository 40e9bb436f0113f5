"""On-box federation runtime: one process per GPU, K clients per GPU, fused in-kernel aggregation.

This is the B200-native replacement of the reference's whole server/client exchange for clients that
live on the same 8-GPU box:

===========================================  =======================================================
reference                                     here
===========================================  =======================================================
N gRPC uploads of pickled weights             each rank publishes ONE partial (weighted sum of its K
(fl_client.py:61-65 -> fl_server.py:176-179)  local clients, kernel K14) in an IPC-exported arena
NumPy mean on the server CPU                  shard owner pulls the G partials over NVLink with 128-bit
(fl_server.py:92-102)                         loads inside ``crackfl_fedagg`` and updates its fp32 master shard
N unicast downloads (fl_server.py:190,201)    the same kernel stores the new shard into every rank's
                                              global arena (peer stores / multimem.st)
count-to-N + 20 s polling (fl_client.py:141)  signal-pad barrier inside the kernel, with a spin timeout
===========================================  =======================================================

``torch.distributed`` (NCCL or Gloo) is used only to exchange the 64-byte IPC handles at start-up and
for the max-over-ranks of timings.  On CPU (tests, Gloo) the aggregation falls back to an all_reduce
of the flat arenas with the same API, so the host-side logic is exercised without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from ..io import weights as wio
from ..models.spec import REFERENCE_SPEC, UNetSpec


class RawCudaBuffer:
    """cudaMalloc'd buffer (own allocation => IPC-exportable base pointer) viewable as a torch tensor."""

    def __init__(self, lib, nbytes: int, ptr: Optional[int] = None, owner: bool = True):
        self.lib, self.nbytes, self.owner = lib, nbytes, owner
        if ptr is None:
            out = C.c_void_p()
            rc = lib.crackfl_dev_malloc(C.byref(out), nbytes)
            if rc:
                raise RuntimeError(f"cudaMalloc({nbytes}) failed: {rc}")
            ptr = out.value
        self.ptr = ptr

    def tensor(self, dtype: torch.dtype, numel: int, device) -> torch.Tensor:
        typestr = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.uint32: "<u4", torch.int32: "<i4",
                   torch.uint8: "|u1"}[dtype]

        class _View:
            pass
        v = _View()
        v.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (self.ptr, False), "version": 3}
        t = torch.as_tensor(v, device=device)
        if dtype == torch.bfloat16:
            t = t.view(torch.bfloat16)
        t._crackfl_keepalive = self
        return t

    def export(self) -> bytes:
        h = C.create_string_buffer(64)
        rc = self.lib.crackfl_ipc_export(self.ptr, h)
        if rc:
            raise RuntimeError(f"cudaIpcGetMemHandle failed: {rc}")
        return h.raw


class OnBoxFederation:
    """Per-rank handle of a federation whose clients all live on this box.

    ``payload``: "fp32" (exact reference mean) or "bf16_delta" (partials travel as bf16 deltas against the
    global model: half the NVLink bytes; the fp32 master keeps the accumulation exact to bf16 rounding
    of the *update*, not of the weights).

    Memory (CUDA):  ``global_arena`` is an IPC-exported buffer.  With one client per GPU it is meant to BE the
    engine's parameter arena (``UNetEngine(..., arena=fed.global_arena)``): peers pull this rank's trained weights
    straight out of it and the shard owners store the new global model straight into it, so the broadcast *is* the
    weight reload (K13) and a round boundary moves no byte on this GPU beyond what the kernel itself moves.  That is
    safe in place because shard r of every arena is read only by rank r, which reads element i of all arenas before
    it writes element i.  With K > 1 clients per GPU the clients share one engine, so each trained arena is copied
    into ``updates[k]`` (device to device), pre-reduced (K14) into the exported ``partial`` and the engine reloads from
    ``global_arena`` at the start of each fit.

    Failure handling: spin waits inside the kernel time out (``timeout_s``, default 60 s) and raise a flag; the flag
    is copied to pinned host memory behind the launch and looked at when the NEXT call starts (or in ``check()``), so
    a healthy round never synchronises the stream.  After a time-out the flag and the counters are reset and the
    federation can go on.  ``rendezvous()`` is an optional host-side barrier (Gloo, no GPU work) for callers whose
    ranks may reach the boundary many seconds apart.
    """

    def __init__(self, spec: UNetSpec = REFERENCE_SPEC, clients_per_gpu: int = 1, strategy: str = "fedavg",
                 payload: str = "fp32", eta=1e-2, beta1=0.9, beta2=0.99, tau=1e-3, numel: Optional[int] = None,
                 grid_blocks: int = 148 * 8, nvls: Optional[bool] = None, timeout_s: float = 60.0,
                 remote_slots: int = 0):
        self.spec = spec
        self.n = int(numel if numel is not None else spec.arena_numel)
        self.mask_from_spec = numel is None  # synthetic sizes (bandwidth sweeps) treat every element as trainable
        if self.n % 4:
            raise ValueError("arena length must be a multiple of 4")
        self.K = clients_per_gpu
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.strategy, self.payload = strategy, payload
        self.server_opt = 1 if strategy == "fedadam" else 0
        self.hp = (eta, beta1, beta2, tau)
        self.cuda = torch.cuda.is_available()
        self.epoch = 0
        self.nvls = bool(int(os.environ.get("CRACKFL_NVLS", "0"))) if nvls is None else nvls
        self.mc = {}
        self.grid_blocks = int(os.environ.get("CRACKFL_AGG_GRID", grid_blocks))
        self.timeout_s = float(os.environ.get("CRACKFL_AGG_TIMEOUT_S", timeout_s))
        self.remote_slots = int(remote_slots)  # staging slots for off-box (gRPC) updates; only rank 0 fills them
        self._remote: List[float] = []  # sample counts of the remote updates staged for the coming round
        self._events = None
        self._err_pending = None
        self._host_group = None
        self._last_ms = 0.0
        if self.cuda:
            self._init_cuda()
        else:
            self._init_cpu()

    # ------------------------------------------------------------------ set-up
    def _init_cpu(self):
        n, K = self.n, self.K
        self.updates = torch.zeros((K + self.remote_slots, n), dtype=torch.float32)
        self.global_arena = torch.zeros(n, dtype=torch.float32)
        self.master = torch.zeros(n, dtype=torch.float32)
        self.m = torch.zeros(n, dtype=torch.float32)
        self.v = torch.zeros(n, dtype=torch.float32)
        self.trainable = torch.ones(n, dtype=torch.bool)
        if self.mask_from_spec:
            for t in self.spec.tensors:
                self.trainable[t.offset:t.offset + t.numel] = t.trainable

    def _init_cuda(self):
        from ..ops import lib as L
        self.L, self.lib = L, L.get_lib()
        n, K, world = self.n, self.K, self.world
        self.dev = torch.device("cuda", torch.cuda.current_device())
        pbytes = 2 if self.payload == "bf16_delta" else 4
        self.nvls = self.nvls and world > 1
        self._raw = {"pads": RawCudaBuffer(self.lib, 2 * L.MAX_RANKS * 4)}
        self._symm = {}
        if self.nvls:
            # NVLS: arenas come from torch's symmetric-memory allocator, which also maps them on an NVSwitch
            # multicast address (multimem.ld_reduce for the pull, multimem.st for the broadcast)
            import torch.distributed._symmetric_memory as symm
            self.partial = symm.empty(n, dtype=torch.bfloat16 if pbytes == 2 else torch.float32, device=self.dev)
            self.global_arena = symm.empty(n, dtype=torch.float32, device=self.dev)
            self.partial.zero_(); self.global_arena.zero_()
            for name, t in (("partial", self.partial), ("global", self.global_arena)):
                h = symm.rendezvous(t, group=dist.group.WORLD)
                if not getattr(h, "multicast_ptr", 0):
                    raise RuntimeError("symmetric memory has no multicast mapping on this system (NVLS unavailable)")
                self._symm[name] = h
                self.mc[name] = int(h.multicast_ptr)
        else:
            self._raw["partial"] = RawCudaBuffer(self.lib, n * pbytes)
            self._raw["global"] = RawCudaBuffer(self.lib, n * 4)
            self.partial = self._raw["partial"].tensor(torch.bfloat16 if pbytes == 2 else torch.float32, n, self.dev)
            self.global_arena = self._raw["global"].tensor(torch.float32, n, self.dev)
        # K == 1: the single client's update is read where it was trained (the live arena); no updates[] slot for it
        self.live = (K == 1)
        nslots = (0 if self.live else K) + self.remote_slots
        self.updates = torch.zeros((max(nslots, 1), n), dtype=torch.float32, device=self.dev)
        self.master = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.m = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.v = torch.zeros(n, dtype=torch.float32, device=self.dev)
        mask = np.ones(n, np.uint8)
        if self.mask_from_spec:
            for t in self.spec.tensors:
                mask[t.offset:t.offset + t.numel] = 1 if t.trainable else 0
        self.trainable = torch.from_numpy(mask).to(self.dev)
        self.local_sync = torch.zeros(2, dtype=torch.int32, device=self.dev)
        self.error_flag = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self._err_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._stage_host = None
        # peer mapping
        self.peer: Dict[str, List[int]] = {k: [0] * world for k in ("partial", "global", "pads")}
        for name, h in self._symm.items():
            self.peer[name] = [int(p) for p in h.buffer_ptrs]
        if world > 1:
            handles = {k: b.export() for k, b in self._raw.items()}
            gathered: List[Optional[dict]] = [None] * world
            dist.all_gather_object(gathered, {"dev": torch.cuda.current_device(), "h": handles})
            for r, info in enumerate(gathered):
                for k in self._raw:
                    if r == self.rank:
                        self.peer[k][r] = self._raw[k].ptr
                    else:
                        out = C.c_void_p()
                        rc = self.lib.crackfl_ipc_open(info["h"][k], C.byref(out))
                        if rc:
                            raise RuntimeError(f"cudaIpcOpenMemHandle(rank {r}, {k}) failed: {rc}")
                        self.peer[k][r] = out.value
            dist.barrier()
        else:
            for k in self._raw:
                self.peer[k][0] = self._raw[k].ptr
        # shard of this rank (multiple of 4 elements)
        n4 = n // 4
        per = (n4 + world - 1) // world
        self.begin = min(n4, per * self.rank) * 4
        self.end = min(n4, per * (self.rank + 1)) * 4

    # ------------------------------------------------------------------ API
    def set_global(self, flat: torch.Tensor):
        """Install the initial global model on this rank (same values on every rank)."""
        self.global_arena.copy_(flat.to(self.global_arena.device))
        self.master.copy_(self.global_arena)

    def publish(self, k: int, flat: torch.Tensor):
        """Client k (local index) hands in its trained weights.  A no-op when ``flat`` is the exported arena itself
        (one client per GPU training in place); otherwise one device-to-device copy into the client's slot."""
        if self.cuda and self.live:
            if flat.data_ptr() != self.global_arena.data_ptr():
                self.global_arena.copy_(flat)
            return
        self.updates[k].copy_(flat)

    def stage_remote(self, flat_host, num_samples: float):
        """(K16) Rank 0 only: stage one off-box client's update (a flat fp32 host array, e.g. a decoded gRPC 'D'
        payload, fl_server.py:176-179) into the next free remote slot; it joins the coming round as one more client with
        its own sample count.  Host -> pinned -> device, asynchronous on the current stream."""
        if self.rank != 0:
            raise RuntimeError("remote updates are staged on rank 0 (where the gRPC service terminates)")
        j = len(self._remote)
        if j >= self.remote_slots:
            raise ValueError(f"{self.remote_slots} remote slots, all in use")
        src = torch.as_tensor(np.ascontiguousarray(flat_host, dtype=np.float32)).reshape(-1)
        if src.numel() != self.n:
            raise ValueError(f"remote update has {src.numel()} elements, expected {self.n}")
        slot = (0 if (self.cuda and self.live) else self.K) + j
        if self.cuda:
            if self._stage_host is None:  # one pinned buffer per slot: no wait between consecutive uploads
                self._stage_host = torch.empty((self.remote_slots, self.n), dtype=torch.float32).pin_memory()
            self._stage_host[j].copy_(src)
            self.updates[slot].copy_(self._stage_host[j], non_blocking=True)
        else:
            self.updates[slot].copy_(src)
        self._remote.append(float(num_samples))

    def configure(self, num_samples_local: Sequence[float], weighted: bool = True):
        """Control plane, once per federation (or whenever sample counts / the set of remote updates change):
        exchange the per-client sample counts and derive every client's coefficient n_k / sum n.  This is the only
        collective call; ``aggregate`` itself makes none."""
        ns = [float(x) if weighted else 1.0 for x in num_samples_local]
        if len(ns) != self.K:
            raise ValueError(f"need {self.K} sample counts, got {len(ns)}")
        rem = [float(x) if weighted else 1.0 for x in self._remote]
        mine = {"ns": ns, "remote": rem}
        allinfo: List[Optional[dict]] = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(allinfo, mine)
        else:
            allinfo = [mine]
        tot = float(sum(sum(r["ns"]) + sum(r["remote"]) for r in allinfo))
        if tot <= 0:
            raise ValueError("total sample count must be positive")
        self.coef_all = [[x / tot for x in r["ns"]] for r in allinfo]
        self.coef_remote = [[x / tot for x in r["remote"]] for r in allinfo]
        self._configured = (tuple(ns), weighted, tuple(rem))

    def rendezvous(self):
        """Host-side barrier (Gloo over the store; no GPU work, no NCCL): ranks whose local fits end seconds apart
        wait here on the CPU instead of inside the kernel's spin loop."""
        if self.world > 1:
            if self._host_group is None:
                self._host_group = dist.new_group(backend="gloo")
            dist.barrier(group=self._host_group)

    def align(self):
        """Device-side rendezvous: a barrier-only launch of the aggregation kernel (empty shard).  Kernels enqueued
        behind it start within microseconds of each other on all ranks; benchmarks use it so that the time they report
        for ``aggregate`` is the kernel and not the skew with which the ranks' hosts reached the launch."""
        if not self.cuda or self.world == 1:
            return
        L = self.L
        self.epoch += 1
        q = L.FedAggParams()
        for r in range(self.world):
            q.client[r] = self.peer["global"][r]
            q.out[r] = self.peer["global"][r]
            q.pads[r] = self.peer["pads"][r]
        q.num_clients = q.num_out = self.world
        q.begin = q.end = self.begin
        q.rank, q.world, q.epoch = self.rank, self.world, self.epoch
        q.error_flag, q.local_sync = self.error_flag.data_ptr(), self.local_sync.data_ptr()
        q.timeout_cycles = int(self.timeout_s * 1.9e9)
        L.check(self.lib.crackfl_fedagg(C.byref(q), 1, torch.cuda.current_stream(self.dev).cuda_stream), "fedagg align")

    def check(self):
        """Raise if the previous aggregation timed out at a barrier (and reset, so the next round can run)."""
        if not self.cuda or self._err_pending is None:
            return
        self._err_pending.synchronize()
        self._err_pending = None
        flag = int(self._err_host[0])
        if flag != 0:
            self.error_flag.zero_(); self.local_sync[1:].zero_()
            self._err_host.zero_()
            raise RuntimeError(f"rank {self.rank}: aggregation barrier timed out (phase flag {flag}); state was reset")

    @property
    def last_agg_ms(self) -> float:
        """Device time of the most recent aggregation (CUDA events around its launches; waits for it to finish)."""
        if self._events is not None:
            e0, e1 = self._events
            e1.synchronize()
            self._last_ms = e0.elapsed_time(e1)
            self._events = None
        return self._last_ms

    @last_agg_ms.setter
    def last_agg_ms(self, v: float):
        self._last_ms, self._events = float(v), None

    def aggregate(self, num_samples_local: Optional[Sequence[float]] = None, weighted: bool = True,
                  reduce_timing: bool = False) -> Optional[float]:
        """One FedAvg/FedAdam aggregation + broadcast, enqueued on the current stream: no host synchronisation, no
        collective call.  With ``reduce_timing`` it waits and returns the device time in ms, max over ranks (that
        all_reduce is for reporting only); otherwise read ``last_agg_ms`` when the number is wanted."""
        if num_samples_local is not None:
            key = (tuple(float(x) if weighted else 1.0 for x in num_samples_local), weighted,
                   tuple(float(x) if weighted else 1.0 for x in self._remote))
            # with remote slots every rank re-runs the exchange every round: only rank 0 knows how many off-box
            # updates joined, and a collective that some ranks skip would hang
            if self.remote_slots > 0 or getattr(self, "_configured", None) != key:
                self.configure(num_samples_local, weighted)
        if getattr(self, "coef_all", None) is None:
            raise RuntimeError("call configure(num_samples) first")
        coef = torch.tensor(self.coef_all[self.rank], dtype=torch.float32)
        rcoef = list(self.coef_remote[self.rank])
        self.epoch += 1
        if not self.cuda:
            ms = self._aggregate_cpu(coef, rcoef)
            self._remote = []
            return ms
        self.check()
        L, lib = self.L, self.lib
        st = torch.cuda.current_stream(self.dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        K = self.K
        p = None
        nrem = len(rcoef)
        # which buffer does each rank offer to its peers?  The live arena itself when that rank's single local update
        # needs no pass of its own (no remote updates, fp32 payload, no NVLS pre-scaling), else its pre-reduced partial.
        def is_direct(r):
            return K == 1 and self.payload == "fp32" and not self.nvls and len(self.coef_remote[r]) == 0
        if not is_direct(self.rank):
            # (K14) local pre-reduce of this GPU's K clients (+ staged remote updates) -> partial
            p = L.FedAggParams()
            srcs = [self.global_arena.data_ptr()] if self.live else [self.updates[k].data_ptr() for k in range(K)]
            base = 0 if self.live else K
            srcs += [self.updates[base + j].data_ptr() for j in range(nrem)]
            cs = [float(c) for c in coef] + [float(c) for c in rcoef]
            for i, (ptr, c) in enumerate(zip(srcs, cs)):
                p.client[i] = ptr
                p.coef[i] = c
            nc = len(srcs)
            if self.payload == "bf16_delta":
                p.client[nc] = self.master.data_ptr()  # partial = sum c_k (w_k - w_global); master holds w_global
                p.coef[nc] = -float(sum(cs))
                nc += 1
            p.num_clients, p.num_out = nc, 1
            p.out[0] = self.partial.data_ptr()
            p.payload_bf16, p.out_bf16 = 0, 1 if self.payload == "bf16_delta" else 0
            p.delta = p.server_opt = 0
            p.master = p.m = p.v = p.trainable = p.mc_out = 0
            p.begin, p.end = 0, self.n
            p.rank, p.world, p.epoch = 0, 1, 1
            p.error_flag = p.local_sync = 0
            p.load_mode = 1
        # (K11-K13) cross-GPU reduce + server update + broadcast, one launch
        q = L.FedAggParams()
        for r in range(self.world):
            d = is_direct(r)
            q.client[r] = self.peer["global"][r] if d else self.peer["partial"][r]
            q.coef[r] = float(self.coef_all[r][0]) if d else 1.0
            q.out[r] = self.peer["global"][r]
            q.pads[r] = self.peer["pads"][r]
        q.num_clients = q.num_out = self.world
        q.payload_bf16 = 1 if self.payload == "bf16_delta" else 0
        q.out_bf16 = 0
        q.delta = 1 if self.payload == "bf16_delta" else 0
        q.server_opt = self.server_opt
        q.eta, q.beta1, q.beta2, q.tau = self.hp
        q.master, q.m, q.v = self.master.data_ptr(), self.m.data_ptr(), self.v.data_ptr()
        q.trainable = self.trainable.data_ptr()
        q.mc_out = self.mc.get("global", 0)
        q.mc_in = self.mc.get("partial", 0)
        q.begin, q.end = self.begin, self.end
        q.rank, q.world, q.epoch = self.rank, self.world, self.epoch
        q.error_flag, q.local_sync = self.error_flag.data_ptr(), self.local_sync.data_ptr()
        q.load_mode = int(os.environ.get("CRACKFL_AGG_LOAD_MODE", "0"))
        q.unroll = int(os.environ.get("CRACKFL_AGG_UNROLL", "4"))
        q.timeout_cycles = int(self.timeout_s * 1.9e9)
        # (the parameter blocks are built before the first event so that, on an idle stream, the events bracket the
        # launches and not the host-side set-up)
        e0.record(st)
        if p is not None:
            L.check(lib.crackfl_fedagg(C.byref(p), 148 * 8, st.cuda_stream), "fedagg pre-reduce")
        L.check(lib.crackfl_fedagg(C.byref(q), self.grid_blocks, st.cuda_stream), "fedagg")
        e1.record(st)
        self._events = (e0, e1)
        if self.world > 1:  # time-out flag: copied behind the launch, looked at when the next call starts
            self._err_host.copy_(self.error_flag, non_blocking=True)
            self._err_pending = torch.cuda.Event()
            self._err_pending.record(st)
        self._remote = []
        if not reduce_timing:
            return None
        ms = torch.tensor([self.last_agg_ms], device=self.dev)
        self.check()
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        self.last_agg_ms = float(ms.item())
        return self._last_ms

    def _aggregate_cpu(self, coef: torch.Tensor, rcoef: Sequence[float]) -> float:
        import time
        t0 = time.perf_counter()
        acc = (self.updates[:self.K] * coef[:, None]).sum(0)
        for j, c in enumerate(rcoef):
            acc = acc + float(c) * self.updates[self.K + j]
        if self.world > 1:
            dist.all_reduce(acc)
        if self.server_opt:
            eta, b1, b2, tau = self.hp
            d = acc - self.master
            tr = self.trainable
            self.m[tr] = b1 * self.m[tr] + (1 - b1) * d[tr]
            self.v[tr] = b2 * self.v[tr] + (1 - b2) * d[tr] * d[tr]
            new = acc.clone()
            new[tr] = self.master[tr] + eta * self.m[tr] / (self.v[tr].sqrt() + tau)
            acc = new
        self.master.copy_(acc)
        self.global_arena.copy_(acc)
        self.last_agg_ms = (time.perf_counter() - t0) * 1e3
        return self._last_ms

    def global_weights(self) -> List[np.ndarray]:
        return wio.unflatten(self.global_arena.detach().cpu().numpy(), self.spec)


def init_distributed(backend: Optional[str] = None):
    """Initialise torch.distributed from the torchrun environment (no-op for a single process)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1 or dist.is_initialized():
        return
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if torch.cuda.is_available():
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group(backend=backend)


def shutdown_distributed():
    """Counterpart of init_distributed for CLI exits: ranks leave together and the process group is torn down."""
    if dist.is_available() and dist.is_initialized():
        try:
            dist.barrier()
        finally:
            dist.destroy_process_group()


def run_federated_onbox(cfg, rounds: int, clients_per_gpu: int = 1, strategy: str = "fedavg", payload: str = "fp32",
                        spec: Optional[UNetSpec] = None, steps_per_round: Optional[int] = None, log=print,
                        eval_global: bool = False, bridge=None, remote_slots: int = 0) -> dict:
    """Full on-box federation: every rank hosts ``clients_per_gpu`` clients that fit locally and aggregate in-kernel
    (the round loop of fl_client.py:121-166 + fl_server.py:107-135 without the wire).

    On CUDA nothing leaves the device between rounds: with one client per GPU the engine trains IN the exported
    arena, so the fused kernel pulls the trained weights from where they are and its broadcast is the reload; with
    K > 1 clients the shared engine reloads from / publishes to device buffers (two 8 MB device copies per client).
    ``bridge`` (``parallel.bridge.OnBoxBridge``, rank 0; every rank passes the same ``remote_slots``): off-box gRPC
    clients served by an ``FLServer`` whose aggregator is the bridge join each round as extra clients.
    ``eval_global``: score each new global model on the first client's validation set (inference-mode BN).

    Returns {"rounds": [...per-round records...], "weights": final global weights (rank 0)}.
    """
    from ..trainer import LocalTrainer, make_backend, spec_from_config
    from ..data.generator import make_datasets
    spec = spec or spec_from_config(cfg)
    fed = OnBoxFederation(spec, clients_per_gpu, strategy, payload, cfg.server.fedadam_eta, cfg.server.fedadam_beta1,
                          cfg.server.fedadam_beta2, cfg.server.fedadam_tau,
                          remote_slots=remote_slots)
    live = fed.cuda and clients_per_gpu == 1
    backend = make_backend(cfg, spec, arena=fed.global_arena if live else None)
    device_path = fed.cuda and backend.name == "b200"
    init = torch.from_numpy(wio.flatten(spec.init_weights(cfg.server.seed), spec))
    fed.set_global(init)
    trainers = []
    for k in range(clients_per_gpu):
        cid = fed.rank * clients_per_gpu + k
        tr, va = make_datasets(cfg, cid)
        trainers.append(LocalTrainer(cfg, tr, va, spec=spec, backend=backend, client_index=cid, log=lambda s: None))
    records = []
    for rnd in range(1, rounds + 1):
        ns, metrics = [], []
        for k, t in enumerate(trainers):
            if device_path:
                eng = backend.engine
                if not live:
                    eng.load_flat(fed.global_arena)  # device-to-device
                _, m = t.fit(None, rnd, return_weights=False)
                fed.publish(k, eng.arena)  # no-op when the engine trains in the exported arena
            else:
                w2, m = t.fit(fed.global_weights(), rnd)
                fed.publish(k, torch.from_numpy(wio.flatten(w2, spec)).to(fed.updates.device))
            ns.append(m["num_samples"]); metrics.append(m)
        if bridge is not None and fed.rank == 0:
            bridge.collect(fed, rnd)
        fed.rendezvous()  # host-side: ranks may finish their fits seconds apart
        fed.aggregate(ns, weighted=(strategy != "fedavg"))
        ms = fed.last_agg_ms
        if bridge is not None and fed.rank == 0:
            bridge.deliver(fed)
        rec = {"round": rnd, "aggregate_ms": ms, "loss": float(np.mean([m["loss"] for m in metrics])),
               "accuracy": float(np.mean([m["accuracy"] for m in metrics])),
               "images_per_sec_per_client": float(np.mean([m["images_per_sec"] for m in metrics])),
               "fit_seconds": float(np.sum([m["fit_seconds"] for m in metrics]))}
        if eval_global and trainers[0].val_ds is not None and len(trainers[0].val_ds) > 0:
            be = backend
            if device_path:
                if not live:
                    be.engine.load_flat(fed.global_arena)
            else:
                be.begin_round(fed.global_weights())
            be.reset_metrics()
            for bi in range(len(trainers[0].val_ds)):
                be.eval_batch(*trainers[0].val_ds[bi])
            ev = be.read_metrics()
            rec["global_val_loss"], rec["global_val_accuracy"] = ev["loss"], ev["accuracy"]
        records.append(rec)
        if fed.rank == 0:
            log(f"[onbox] round {rnd}: loss {rec['loss']:.4f} acc {rec['accuracy']:.4f} aggregate {ms:.3f} ms"
                + (f" global val_loss {rec['global_val_loss']:.4f}" if "global_val_loss" in rec else ""))
    fed.check()
    return {"rounds": records, "weights": fed.global_weights() if fed.rank == 0 else None, "federation": fed}
