"""B200 U-Net engine: the whole train / eval / predict step as a static schedule of hand-written
sm_100a kernels over pre-allocated NHWC bf16 buffers, captured in CUDA graphs.

This is the device counterpart of ``change_model_layers`` + ``compile`` + ``fit`` / ``predict``
(``client_fit_model.py:92-174``).  There is no autograd and no tracing compiler: forward,
backward and the Keras-form Adam update are an explicit list of launches (``self.train_prog``)
replayed as one ``cudaGraph`` per step.

Design notes (what differs from the graph as written, with identical results):

* parameters live in ONE flat fp32 arena in Keras ``get_weights()`` order/layout
  (``models/spec.py``); FedAvg, pickle, h5 and checkpoints are views of it.  A table-driven
  ``weight_prep`` launch re-lays the GEMM-shaped kernels out as bf16 K-major operands each step;
* nearest ``UpSampling2D`` commutes with 1x1 convs, BN-apply, add, ReLU and the sigmoid, so the
  decoder residual conv runs at low resolution and the head (1x1 conv + sigmoid + BCE) runs at
  H/2 against the 2x2 mask pixels each logit covers: the (B,H,W,32) tensor never exists;
* BN(train) needs a grid-wide reduction, so each BN is stats -> finalize -> apply fused into
  its consumer (depthwise load, max-pool tail, decoder combine) instead of a separate pass;
* a conv bias followed by training-mode BN has an exactly-zero gradient; it is not computed.

GEMM-shaped ops (pointwise / residual 1x1 convs, ConvT-as-3x3-conv, their dgrad and wgrad) go
to the tcgen05/TMEM/TMA kernels (``csrc/conv_tc.cu``) when ``conv_impl="tc"``, and to the
CUDA-core cross-check kernels (``csrc/conv_simt.cu``) when ``conv_impl="simt"``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..io import weights as wio
from ..ops import lib as L
from .spec import REFERENCE_SPEC, UNetSpec


def default_conv_impl() -> str:
    return os.environ.get("CRACKFL_CONV_IMPL", "tc")


class UNetEngine:
    def __init__(self, spec: UNetSpec = REFERENCE_SPEC, batch: int = 16, height: int = 128, width: int = 128,
                 device: Optional[int] = None, lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999,
                 eps: float = 1e-7, loss: str = "bce", conv_impl: Optional[str] = None, cuda_graph: bool = True,
                 dice_smooth: float = 1.0, bce_clip: bool = False, arena: Optional[torch.Tensor] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("UNetEngine needs a CUDA device (sm_100a)")
        spec.check_input(height, width)
        self.lib = L.get_lib()
        self.spec, self.B, self.H, self.W = spec, batch, height, width
        self.dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self.lr, self.b1, self.b2, self.eps = lr, beta1, beta2, eps
        self.use_dice = 1 if loss == "bce_dice" else 0
        self.dice_smooth = dice_smooth
        # 0: cross-entropy from the logit (what Keras 2 `fit` computes for a sigmoid output layer); 1: clipped probabilities
        self.bce_clip = 1 if bce_clip else 0
        self._ext_arena = arena
        self.conv_impl = conv_impl or default_conv_impl()
        self.cuda_graph = cuda_graph
        self.side_streams = os.environ.get("CRACKFL_SIDE_STREAM", "1") != "0"
        self.launches_per_step = 0
        self._bufs: Dict[str, torch.Tensor] = {}
        self._fins: list = []
        self.fuse_finalize = os.environ.get("CRACKFL_FUSE_FINALIZE", "1") != "0"
        self.dw_tma = os.environ.get("CRACKFL_DW_TMA", "1") != "0"
        self.gemm_stats = os.environ.get("CRACKFL_GEMM_STATS", "1") != "0"
        self.dw_merged = os.environ.get("CRACKFL_DW_MERGED", "1") != "0"
        self.pool_bwd_fused = os.environ.get("CRACKFL_POOL_BWD_FUSED", "1") != "0"
        self.dw_bn_fused = os.environ.get("CRACKFL_DW_BN_FUSED", "1") != "0"
        # measured: batching the split reductions per block (7 launches instead of 22) is 35-45 us SLOWER - the partial
        # tiles fall out of L2 before they are reduced - so the per-layer reduction right behind each GEMM stays the default
        self.wgrad_deferred = os.environ.get("CRACKFL_WGRAD_DEFERRED", "0") != "0"
        self.wgrad_split = os.environ.get("CRACKFL_WGRAD_SPLIT", "1") != "0"  # measured: -20 us per step
        # K1: BN-apply + ReLU inside the second decoder conv of each block (no bn_act pass); depthwise forward on the tensor
        # cores through block-diagonal weight tiles (csrc/conv_tc.cu conv_tc_xf_k)
        # (measured, bench/xf_bench.py + full step: the fused prologue wins 2-4 us per layer for the 64..256-channel decoder
        # convs and loses 10 us on the 32-channel one at 128^2, whose 64-byte halo rows are TMA-bound; the tensor-core
        # depthwise kernel has a floor of ~2 us per 128-pixel tile per SM WITHOUT any transform work - the 180 x 128 B
        # halo load and the 128 x 128 B store go through the one TMA unit of the SM - which is what the CUDA-core stencil
        # already reaches, so it stays off: CRACKFL_XF_DEC=2 forces the prologue for every layer, CRACKFL_XF_DW=1 the
        # depthwise path)
        # K2: BN-backward apply inside the decoder convs' input-gradient kernels (no bn_bwd_apply pass; 1 = blocks with
        # f >= 64, 2 = all).  Measured on the full step: 1.160 ms with it, 1.116 without (two halo tiles per stage through
        # six transform warps cost more than the bn_bwd_apply pass they remove), so it is off unless asked for.
        self.xf_bwd = int(os.environ.get("CRACKFL_XF_BWD", "0"))
        # K1: BN-apply + ReLU as the prologue of the decoder's second conv.  1.122 ms with it, 1.115 without: off by default.
        self.xf_dec = int(os.environ.get("CRACKFL_XF_DEC", "0"))
        self.adam_early = os.environ.get("CRACKFL_ADAM_EARLY", "1") != "0"
        # decoder: down2sum + BN2-backward reductions + next down2sum in one launch (5 launches fewer).  Measured on the full
        # step: 1.024 ms with it, 1.023 without - off by default, kept as a tested entry point
        self.d2s_bn = os.environ.get("CRACKFL_D2S_BN", "0") != "0"
        self.prep_flat = os.environ.get("CRACKFL_PREP_FLAT", "1") != "0"  # one 32x32 tile per block (0: 48 blocks per operand)
        self.xf_dw = os.environ.get("CRACKFL_XF_DW", "0") != "0"
        # partial tiles added straight into the gradient arena with vector reductions: no workspace, no reduction launch
        # (22 launches fewer per step, 1.263 vs 1.287 ms/step; CRACKFL_WGRAD_ATOMIC=0 restores workspace + reduction)
        self.wgrad_atomic = os.environ.get("CRACKFL_WGRAD_ATOMIC", "1") != "0"
        self._wg_descs, self._wg_off, self._wg_max, self._wg_flushed = [], 0, 0, 0
        with torch.cuda.device(self.dev):
            self._alloc_params()
            self._build_ops_table()
            self._alloc_activations()
            self.train_prog = self._build_train()
            self.eval_prog = self._build_forward(training=False, with_mask=True, prob=False)
            self.pred_prog = self._build_forward(training=False, with_mask=False, prob=True)
        skip = [k for k in os.environ.get("CRACKFL_SKIP_KERNELS", "").split(",") if k]
        if skip:  # profiling aid only ("what if this kernel were free"): the step no longer trains correctly
            self.train_prog = [e for e in self.train_prog if e[0] is None or not any(k in e[0].__name__ for k in skip)]
        self._graphs: Dict[str, torch.cuda.CUDAGraph] = {}
        # stream priorities (captured into the graph's kernel nodes): the main chain is the critical path of the step, the
        # weight-gradient work on the side streams only has to be done before Adam - when both have blocks pending, the
        # main chain's go first.  CRACKFL_PRIO=0 puts everything back on default-priority streams.
        prio = os.environ.get("CRACKFL_PRIO", "0") != "0"  # measured: 1.395 vs 1.292 ms/step - starving the side streams only moves their work to the end
        lo = 0
        self._main_prio = -1 if prio else 0
        self._side_stream = torch.cuda.Stream(self.dev, priority=lo)
        self._aux_stream = torch.cuda.Stream(self.dev, priority=self._main_prio)
        self._side2_stream = torch.cuda.Stream(self.dev, priority=lo)
        self._side3_stream = torch.cuda.Stream(self.dev, priority=lo)  # bias-gradient column sums
        self.launches_per_step = sum(1 for e in self.train_prog if e[0] is not None)
        # forward + backward without the optimiser launches (gradient checks): the gradient arena is left as computed
        self.grad_prog = [e for e in self.train_prog if e[0] is None or
                          getattr(e[0], "__name__", "") not in ("crackfl_adam_part", "crackfl_adam_step_tail", "crackfl_adam_step")]

    # ------------------------------------------------------------------ allocation
    def _new(self, name: str, shape, dtype=torch.bfloat16) -> torch.Tensor:
        t = torch.zeros(tuple(int(s) for s in shape), dtype=dtype, device=self.dev)
        self._bufs[name] = t
        return t

    def _alloc_params(self):
        n = self.spec.arena_numel
        if self._ext_arena is not None:
            # parameters live in a buffer the caller owns (the on-box runtime's IPC-exported arena: peers read the
            # trained weights out of it and write the next global model into it, runtime.py)
            a = self._ext_arena
            if a.dtype != torch.float32 or a.numel() != n or not a.is_contiguous() or a.device != self.dev:
                raise ValueError("external arena must be a contiguous fp32 tensor of spec.arena_numel elements on the engine's device")
            self.arena = a
            self._bufs["arena"] = a
        else:
            self.arena = self._new("arena", (n,), torch.float32)
        self.grads = self._new("grads", (n,), torch.float32)
        self.m = self._new("adam_m", (n,), torch.float32)
        self.v = self._new("adam_v", (n,), torch.float32)
        self.step = self._new("adam_step", (1,), torch.int32)
        # per-BN scratch: zeroed-every-step region [stats 2C | sums 2C]*, persistent region [bn 4C]*
        self.bn_layers = [l for l in self.spec.layers if l.op == "bn"]
        zoff, poff = 0, 0
        self.bn_z, self.bn_p = {}, {}
        for l in self.bn_layers:
            self.bn_z[l.name] = zoff
            zoff += 4 * l.cout
            self.bn_p[l.name] = poff
            poff += 4 * l.cout
        self.zero_region = self._new("zero_region", (zoff + 16,), torch.float32)
        self.acc_off = zoff  # 6 floats of loss/metric accumulators for the current step
        self.bn_params = self._new("bn_params", (poff,), torch.float32)
        self.metrics_total = self._new("metrics_total", (8,), torch.float32)
        # loss / metric sums of the most recent train step (the Adam launch saves them here before it clears the region)
        self.last_acc = self._new("last_acc", (8,), torch.float32)
        self._region_dirty = False  # set by anything but a complete train step; the next train step then clears it first

    def P(self, tensor_idx: int) -> int:
        """Device address of weight tensor ``tensor_idx`` inside the arena."""
        return self.arena.data_ptr() + 4 * self.spec.tensors[tensor_idx].offset

    def G(self, tensor_idx: int) -> int:
        return self.grads.data_ptr() + 4 * self.spec.tensors[tensor_idx].offset

    def _build_ops_table(self):
        """bf16 GEMM operands: Wf (forward) and Wd (dgrad) for every GEMM-shaped layer."""
        descs: List[L.PrepDesc] = []
        self.ops: Dict[str, Tuple[int, int, int, int]] = {}  # name -> (elem offset, N, T, K)
        off = 0

        def add(name, src_off, N, T, K, s_n, s_t, s_k, kpad=0):
            nonlocal off
            d = L.PrepDesc(src_off=src_off, dst_off=off, N=N, T=T, K=K, pad_=kpad, s_n=s_n, s_t=s_t, s_k=s_k)
            descs.append(d)
            Kd = kpad or K
            self.ops[name] = (off, N, T, Kd)
            off += (N * T * Kd + 63) // 64 * 64

        # entry conv as a K=32 GEMM over the im2col matrix (27 real rows of the (3,3,3,c0) kernel, 5 zero columns)
        self.entry_tc = (self.conv_impl == "tc" and self.spec.entry % 32 == 0
                         and os.environ.get("CRACKFL_ENTRY_TC", "1") != "0")
        if self.entry_tc:
            c0 = self.spec.entry
            k0 = self.spec.tensors[self.spec.layers[0].tensors[0]]
            add("conv0.f", k0.offset, c0, 1, 27, 1, 0, c0, kpad=32)
        for l in self.spec.layers:
            ts = self.spec.tensors
            if l.op == "sepconv":
                pw = ts[l.tensors[1]]  # (1,1,in,out)
                add(l.name + ".f", pw.offset, l.cout, 1, l.cin, 1, 0, l.cout)
                add(l.name + ".d", pw.offset, l.cin, 1, l.cout, l.cout, 0, 1)
            elif l.op == "conv" and l.ksize == 1 and l.cout >= 8:
                k = ts[l.tensors[0]]
                add(l.name + ".f", k.offset, l.cout, 1, l.cin, 1, 0, l.cout)
                add(l.name + ".d", k.offset, l.cin, 1, l.cout, l.cout, 0, 1)
            elif l.op == "convT":
                k = ts[l.tensors[0]]  # (3,3,out,in); conv-equivalent Wc[t][i][o] = K[8-t][o][i]
                oi = l.cout * l.cin
                add(l.name + ".f", k.offset + 8 * oi, l.cout, 9, l.cin, l.cin, -oi, 1)
                add(l.name + ".d", k.offset, l.cin, 9, l.cout, 1, oi, l.cin)
        # depthwise kernels as block-diagonal bf16 tiles (only the diagonals are rewritten each step)
        self.dw_ops: Dict[str, int] = {}
        ddescs: List[L.DwDiagDesc] = []
        if self.conv_impl == "tc" and self.xf_dw:
            for l in self.spec.layers:
                if l.op == "sepconv" and l.cin % 32 == 0:
                    kc = 64 if l.cin % 64 == 0 else 32
                    dwt = self.spec.tensors[l.tensors[0]]  # (3,3,cin,1)
                    ddescs.append(L.DwDiagDesc(src_off=dwt.offset, dst_off=off, C=l.cin, kc=kc, flip=0, pad_=0))
                    self.dw_ops[l.name] = off
                    off += (l.cin // kc) * 9 * kc * kc
        self.ops_buf = self._new("ops", (off,), torch.bfloat16)
        self.num_ddesc = len(ddescs)
        if ddescs:
            darr = (L.DwDiagDesc * len(ddescs))(*ddescs)
            self.ddesc_dev = torch.frombuffer(bytearray(bytes(darr)), dtype=torch.uint8).clone().to(self.dev)
            self._bufs["dw_diag_descs"] = self.ddesc_dev
        arr = (L.PrepDesc * len(descs))(*descs)
        raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).clone()
        self.desc_dev = raw.to(self.dev)
        self._bufs["prep_descs"] = self.desc_dev
        self.num_desc = len(descs)
        pref = [0]
        for d in descs:
            pref.append(pref[-1] + ((d.N + 31) // 32) * ((d.K + 31) // 32) * d.T)
        self.prep_tiles = pref[-1]
        self.prep_prefix = torch.tensor(pref, dtype=torch.int32).to(self.dev)
        self._bufs["prep_prefix"] = self.prep_prefix

    def op_ptr(self, name: str) -> int:
        return self.ops_buf.data_ptr() + 2 * self.ops[name][0]

    def _alloc_activations(self):
        s, B, H, W = self.spec, self.B, self.H, self.W
        self.x_u8 = self._new("x_u8", (B, H, W, 3), torch.uint8)
        self.mask_u8 = self._new("mask_u8", (B, H, W), torch.uint8)
        self.prob = self._new("prob", (B, H, W), torch.float32)
        r, c = H // 2, s.entry
        rw = W // 2
        n = self._new
        n("y0", (B, r, rw, c)); n("x0", (B, r, rw, c)); n("x0e", (B, r // 2, rw // 2, c))
        n("dz0", (B, r, rw, c)); n("dy0", (B, r, rw, c))
        if self.entry_tc:
            n("x0col", (B, r, rw, 32))  # im2col of the uint8 image: operand of the entry conv and of its weight gradient
        cin = c
        self.enc = []
        for k, f in enumerate(s.down):
            e = dict(k=k, r=r, rw=rw, cin=cin, f=f)
            p = f"e{k}."
            n(p + "d1", (B, r, rw, cin)); n(p + "y1", (B, r, rw, f)); n(p + "d2", (B, r, rw, f)); n(p + "y2", (B, r, rw, f))
            n(p + "rs", (B, r // 2, rw // 2, f)); n(p + "out", (B, r // 2, rw // 2, f))
            n(p + "arg", (B, r // 2, rw // 2, f), torch.uint8)
            if k + 1 < len(s.down):
                n(p + "oute", (B, r // 4, rw // 4, f))
            # backward
            n(p + "Te", (B, r // 2, rw // 2, cin)); n(p + "dy2", (B, r, rw, f)); n(p + "dd2", (B, r, rw, f))
            n(p + "dz1", (B, r, rw, f)); n(p + "dy1", (B, r, rw, f)); n(p + "dd1", (B, r, rw, cin))
            if k > 0:
                n(p + "dx", (B, r, rw, cin))
            self.enc.append(e)
            r //= 2; rw //= 2; cin = f
        n("a0_0", (B, r, rw, cin))  # relu(x3)
        n("dx_bott", (B, r, rw, cin))
        self.dec = []
        for j, f in enumerate(s.up):
            d = dict(j=j, r=r, rw=rw, cin=cin, f=f)
            p = f"d{j}."
            if j > 0:
                n(p + "a0", (B, r, rw, cin))
            n(p + "y1", (B, r, rw, f)); n(p + "a1", (B, r, rw, f)); n(p + "y2", (B, r, rw, f))
            lr_, lrw = (r, rw) if j == 0 else (r // 2, rw // 2)
            n(p + "rl", (B, lr_, lrw, f)); n(p + "s", (B, r, rw, f))
            # backward
            n(p + "ds", (B, r, rw, f))
            if j > 0:
                n(p + "drl", (B, lr_, lrw, f))
            n(p + "dy2", (B, r, rw, f)); n(p + "da1", (B, r, rw, f)); n(p + "dy1", (B, r, rw, f))
            n(p + "da0", (B, r, rw, cin)); n(p + "T", (B, lr_, lrw, cin))
            self.dec.append(d)
            r *= 2; rw *= 2; cin = f
        self.head_c = cin

    # ------------------------------------------------------------------ schedule construction
    def b(self, name: str) -> int:
        return self._bufs[name].data_ptr()

    def _zptr(self, bn_name: str, which: str) -> int:
        C_ = next(l.cout for l in self.bn_layers if l.name == bn_name)
        base = self.zero_region.data_ptr() + 4 * self.bn_z[bn_name]
        return base if which == "stats" else base + 4 * 2 * C_

    def _bnptr(self, bn_name: str) -> int:
        return self.bn_params.data_ptr() + 4 * self.bn_p[bn_name]

    def _acc_ptr(self) -> int:
        return self.zero_region.data_ptr() + 4 * self.acc_off

    def _gemm(self, prog, X, opname, bias, add, Y, B, H, W, stats=0, tag=None):
        off, N, T, K = self.ops[opname]
        extra = (tag,) if tag else ()
        if not self.gemm_stats:
            stats = 0  # tuning knob: BN statistics by a separate pass over the GEMM output instead of the fused epilogue
        if self.conv_impl == "tc":
            prog.append((self.lib.crackfl_conv_tc_fwd, (X, self.op_ptr(opname), bias, add, Y, stats, B, H, W, K, N, T)) + extra)
            return stats != 0
        prog.append((self.lib.crackfl_conv_simt_fwd, (X, self.op_ptr(opname), bias, add, Y, B, H, W, K, N, T)) + extra)
        return False

    def _wgrad(self, prog, X, dY, dW, B, H, W, cin, cout, taps, s_tap, s_ci, s_co, cin_valid=0):
        if self.conv_impl == "tc" and self.wgrad_atomic:
            prog.append((self.lib.crackfl_wgrad_tc_ws_cv, (X, dY, dW, B, H, W, cin, cout, taps, s_tap, s_ci, s_co, 0, 0, cin_valid),
                         "side"))
        elif self.conv_impl == "tc" and self.wgrad_split:
            # partial tiles on the side stream, their reduction on a THIRD stream behind it, two workspace slots in
            # turn: reduction n overlaps GEMM n+1 (GEMM n+2 waits for reduction n before it reuses the slot)
            d = L.WgradRedDesc()
            L.check(self.lib.crackfl_wgrad_plan(B, H, W, cin, cout, taps, s_tap, s_ci, s_co, cin_valid, C.byref(d)),
                    "crackfl_wgrad_plan")
            n = len(self._wg_descs)
            d.dw_off = (dW - self.grads.data_ptr()) // 4
            self._wg_max = max(self._wg_max, (d.need + 255) // 256 * 256)
            self._wg_descs.append(d)
            if n >= 2:
                prog.append((None, None, ("wait", "side", f"r{n - 2}")))
            prog.append(("WGRAD", (X, dY, B, H, W, cin, cout, taps, n % 2, d.need), "side"))
            prog.append((None, None, ("rec", "side", f"w{n}")))
            prog.append((None, None, ("wait", "side2", f"w{n}")))
            prog.append(("WREDUCE", (n, 1), "side2"))
            prog.append((None, None, ("rec", "side2", f"r{n}")))
        elif self.conv_impl == "tc" and self.wgrad_deferred:
            # split-K partial tiles now, ONE reduction launch per block later (_wgrad_flush); the workspace slice is
            # assigned here, the pointers are patched in once the workspace size is known (_wgrad_resolve)
            d = L.WgradRedDesc()
            L.check(self.lib.crackfl_wgrad_plan(B, H, W, cin, cout, taps, s_tap, s_ci, s_co, cin_valid, C.byref(d)),
                    "crackfl_wgrad_plan")
            d.ws_off = self._wg_off
            d.dw_off = (dW - self.grads.data_ptr()) // 4
            self._wg_off += (d.need + 255) // 256 * 256
            self._wg_max = max(self._wg_max, self._wg_off)
            self._wg_descs.append(d)
            prog.append(("WGRAD", (X, dY, B, H, W, cin, cout, taps, d.ws_off, d.need), "side"))
        elif self.conv_impl == "tc":
            if "wgrad_ws" not in self._bufs:  # split-reduction workspace shared by all (stream-ordered) wgrad launches
                self._new("wgrad_ws", (8 << 20,), torch.float32)
            ws = self._bufs["wgrad_ws"]
            prog.append((self.lib.crackfl_wgrad_tc_ws_cv, (X, dY, dW, B, H, W, cin, cout, taps, s_tap, s_ci, s_co,
                                                           ws.data_ptr(), ws.numel(), cin_valid), "side"))
        else:
            prog.append((self.lib.crackfl_wgrad_simt, (X, dY, dW, B, H, W, cin, cout, taps, s_tap, s_ci, s_co), "side"))

    def _wgrad_flush(self, prog):
        """One reduction launch for the wgrad GEMMs recorded since the previous flush; their workspace is reused after it."""
        n = len(self._wg_descs) - self._wg_flushed
        if n > 0 and not self.wgrad_split:
            prog.append(("WREDUCE", (self._wg_flushed, n), "side"))
            self._wg_flushed = len(self._wg_descs)
            self._wg_off = 0

    def _wgrad_resolve(self, prog):
        if not self._wg_descs:
            return prog
        slot = max(self._wg_max, 1024)
        ws = self._new("wgrad_ws", (slot * (2 if self.wgrad_split else 1),), torch.float32)
        if self.wgrad_split:
            for n, d in enumerate(self._wg_descs):
                d.ws_off = (n % 2) * slot
        arr = (L.WgradRedDesc * len(self._wg_descs))(*self._wg_descs)
        self._bufs["wgrad_descs"] = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).clone().to(self.dev)
        dptr, dsz = self._bufs["wgrad_descs"].data_ptr(), C.sizeof(L.WgradRedDesc)
        out = []
        for e in prog:
            if e[0] == "WGRAD":
                X, dY, B, H, W, cin, cout, taps, off, need = e[1]
                if self.wgrad_split:
                    off *= slot
                out.append((self.lib.crackfl_wgrad_tc_deferred, (X, dY, B, H, W, cin, cout, taps, ws.data_ptr() + 4 * off, need),
                            "side"))
            elif e[0] == "WREDUCE":
                i0, n = e[1]
                out.append((self.lib.crackfl_wgrad_reduce_all, (dptr + i0 * dsz, n, ws.data_ptr(), self.grads.data_ptr()), e[2]))
            else:
                out.append(e)
        return out

    def _dw_bwd(self, prog, g, src, w, bn, resid_t, dz, dwk, B, H, W, C, mode, resid, red_bn=0, red_sums=0):
        """Depthwise backward: input gradient on the main stream, weight gradient on the side stream.  Returns True when
        the BN-backward reductions (red_bn / red_sums) were folded into the launch."""
        if self.dw_tma and self.dw_merged:
            # one pass over g and src: warps 0-7 the input gradient (+ BN reductions), warps 8-15 the weight gradient
            if red_bn and self.dw_bn_fused:
                prog.append((self.lib.crackfl_dw3x3_bwd_tma_bn, (g, src, w, bn, resid_t, dz, dwk, B, H, W, C, mode, resid,
                                                                 red_bn, red_sums)))
                return True
            prog.append((self.lib.crackfl_dw3x3_bwd_tma, (g, src, w, bn, resid_t, dz, dwk, B, H, W, C, mode, resid)))
        elif self.dw_tma:
            prog.append((self.lib.crackfl_dw3x3_tma, (g, w, 0, None, dz, src, bn, resid_t, B, H, W, C, 2, 1, mode, resid)))
            prog.append((self.lib.crackfl_dw3x3_wgrad_tma, (g, src, bn, dwk, B, H, W, C, mode), "side"))
        elif W % 4 == 0 and 128 % (C // 8) == 0:
            prog.append((self.lib.crackfl_dw3x3_dgrad, (g, src, w, bn, resid_t, dz, B, H, W, C, mode, resid)))
            prog.append((self.lib.crackfl_dw3x3_wgrad, (g, src, bn, dwk, B, H, W, C, mode), "side"))
        else:
            prog.append((self.lib.crackfl_dw3x3_bwd, (g, src, w, bn, resid_t, dz, dwk, B, H, W, C, mode, resid)))

    def _layer_iter(self):
        """Yield the spec layers grouped as entry / encoder blocks / decoder blocks / head."""
        ls = self.spec.layers
        i = 0
        entry = (ls[0], ls[1]); i = 2
        encs = []
        for _ in self.spec.down:
            encs.append(ls[i:i + 5]); i += 5
        decs = []
        for _ in self.spec.up:
            decs.append(ls[i:i + 5]); i += 5
        return entry, encs, decs, ls[i]

    def _bn_train(self, prog, bnl, y, P, stats_done=False):
        C_ = bnl.cout
        lib = self.lib
        if not stats_done:
            prog.append((lib.crackfl_bn_stats, (y, self._zptr(bnl.name, "stats"), P, C_)))
        g, bt, rm, rv = bnl.tensors
        prog.append((lib.crackfl_bn_finalize, (self._zptr(bnl.name, "stats"), self.P(g), self.P(bt), self.P(rm), self.P(rv),
                                                self._bnptr(bnl.name), C_, float(P), self.spec.bn_eps, self.spec.bn_momentum)))

    def _fin(self, bnl, P) -> "C.POINTER":
        """ctypes crackfl_bn_fin for BN layer ``bnl`` (kept alive on the engine)."""
        g, bt, rm, rv = bnl.tensors
        f = L.BnFin(stats=self._zptr(bnl.name, "stats"), gamma=self.P(g), beta=self.P(bt), run_mean=self.P(rm),
                    run_var=self.P(rv), bn_out=self._bnptr(bnl.name), count=float(P), eps=self.spec.bn_eps,
                    momentum=self.spec.bn_momentum)
        self._fins.append(f)
        return C.byref(f)

    def _bn_eval_all(self, prog):
        """Inference: scale/shift tables of all BN layers from the moving statistics, one launch."""
        if getattr(self, "_bn_eval_descs", None) is None:
            ts = self.spec.tensors
            ds = []
            for l in self.bn_layers:
                g, bt, rm, rv = l.tensors
                ds.append(L.BnEvalDesc(gamma_off=ts[g].offset, beta_off=ts[bt].offset, mean_off=ts[rm].offset,
                                       var_off=ts[rv].offset, bn_off=self.bn_p[l.name], C=l.cout, pad_=0))
            arr = (L.BnEvalDesc * len(ds))(*ds)
            self._bn_eval_descs = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).clone().to(self.dev)
            self._bufs["bn_eval_descs"] = self._bn_eval_descs
        prog.append((self.lib.crackfl_bn_eval_prepare_all, (self.arena.data_ptr(), self.bn_params.data_ptr(),
                                                            self._bn_eval_descs.data_ptr(), len(self.bn_layers),
                                                            self.spec.bn_eps)))

    def _bn_eval(self, prog, bnl):
        g, bt, rm, rv = bnl.tensors
        prog.append((self.lib.crackfl_bn_eval_prepare, (self.P(g), self.P(bt), self.P(rm), self.P(rv), self._bnptr(bnl.name),
                                                        bnl.cout, self.spec.bn_eps)))

    def _build_forward(self, training: bool, with_mask: bool, prob: bool, prog=None):
        lib, B = self.lib, self.B
        prog = [] if prog is None else prog
        (conv0, bn0), encs, decs, head = self._layer_iter()
        b = self.b

        fuse = training and self.fuse_finalize

        def bn(bnl, y, P, stats_done=False, consumer_finalizes=False):
            """stats (+ finalize unless the consumer kernel folds it in) in training, moving statistics in eval."""
            if training:
                if consumer_finalizes:
                    if not stats_done:
                        prog.append((lib.crackfl_bn_stats, (y, self._zptr(bnl.name, "stats"), P, bnl.cout)))
                else:
                    self._bn_train(prog, bnl, y, P, stats_done)
            # (eval: every layer's table was prepared by the single launch at the head of the schedule)

        if not training:
            self._bn_eval_all(prog)
        # bf16 GEMM operands of this step's weights: next to the im2col / entry kernels on the auxiliary stream
        if self.prep_flat and self.num_desc <= 128:
            prog.append((lib.crackfl_weight_prep_flat, (self.arena.data_ptr(), self.ops_buf.data_ptr(), self.desc_dev.data_ptr(),
                                                        self.prep_prefix.data_ptr(), self.num_desc, self.prep_tiles), "aux"))
        else:
            prog.append((lib.crackfl_weight_prep, (self.arena.data_ptr(), self.ops_buf.data_ptr(), self.desc_dev.data_ptr(),
                                                   self.num_desc), "aux"))
        if self.num_ddesc:
            prog.append((lib.crackfl_dw_diag_prep, (self.arena.data_ptr(), self.ops_buf.data_ptr(), self.ddesc_dev.data_ptr(),
                                                    self.num_ddesc), "aux"))
        # entry
        r, rw, c0 = self.H // 2, self.W // 2, self.spec.entry
        if self.entry_tc:
            prog.append((lib.crackfl_entry_im2col, (b("x_u8"), b("x0col"), B, self.H, self.W)))
            prog.append((None, None, "join_aux"))
            sd = self._gemm(prog, b("x0col"), "conv0.f", self.P(conv0.tensors[1]), 0, b("y0"), B, r, rw,
                            self._zptr(bn0.name, "stats") if training else 0)
        else:
            prog.append((lib.crackfl_entry_conv_fwd, (b("x_u8"), self.P(conv0.tensors[0]), self.P(conv0.tensors[1]), b("y0"),
                                                      B, self.H, self.W, c0)))
            prog.append((None, None, "join_aux"))
            sd = False
        bn(bn0, b("y0"), B * r * rw, sd, consumer_finalizes=fuse)
        if fuse:
            prog.append((lib.crackfl_bn_act_fin, (b("y0"), self._fin(bn0, B * r * rw), b("x0"), b("x0e"), B, r, rw, c0, 1)))
        else:
            prog.append((lib.crackfl_bn_act, (b("y0"), self._bnptr(bn0.name), b("x0"), b("x0e"), B, r, rw, c0, 1)))
        xin, xe = "x0", "x0e"
        for e, (sep1, bn1, sep2, bn2, res) in zip(self.enc, encs):
            p = f"e{e['k']}."
            r, rw, cin, f = e["r"], e["rw"], e["cin"], e["f"]
            Pn = B * r * rw
            # residual 1x1 stride-2 conv: a small GEMM that only depends on the block input; it runs on the auxiliary
            # stream next to the separable-conv chain and is joined before the block tail
            self._gemm(prog, b(xe), res.name + ".f", self.P(res.tensors[1]), 0, b(p + "rs"), B, r // 2, rw // 2, tag="aux")
            dw_tc = self.conv_impl == "tc" and self.xf_dw and r >= 16 and rw >= 8
            if dw_tc and sep1.name in self.dw_ops:
                prog.append((lib.crackfl_conv_tc_xf, (b(xin), self.ops_buf.data_ptr() + 2 * self.dw_ops[sep1.name], 0, 0,
                                                      b(p + "d1"), 0, B, r, rw, cin, cin, 1, 0, None, 0, 1)))
            elif self.dw_tma:
                prog.append((lib.crackfl_dw3x3_tma, (b(xin), self.P(sep1.tensors[0]), 0, None, b(p + "d1"), 0, 0, 0,
                                                     B, r, rw, cin, 0, 0, 0, 0)))
            else:
                prog.append((lib.crackfl_dw3x3_fwd, (b(xin), self.P(sep1.tensors[0]), 0, b(p + "d1"), B, r, rw, cin, 0)))
            sd = self._gemm(prog, b(p + "d1"), sep1.name + ".f", self.P(sep1.tensors[2]), 0, b(p + "y1"), B, r, rw,
                            self._zptr(bn1.name, "stats") if training else 0)
            fuse_dw = fuse and (rw % 4 == 0 or self.dw_tma)
            bn(bn1, b(p + "y1"), Pn, sd, consumer_finalizes=fuse_dw)
            if dw_tc and sep2.name in self.dw_ops:
                prog.append((lib.crackfl_conv_tc_xf, (b(p + "y1"), self.ops_buf.data_ptr() + 2 * self.dw_ops[sep2.name], 0, 0,
                                                      b(p + "d2"), 0, B, r, rw, f, f, 2, self._bnptr(bn1.name),
                                                      self._fin(bn1, Pn) if fuse_dw else None, 0, 1)))
            elif self.dw_tma:
                prog.append((lib.crackfl_dw3x3_tma, (b(p + "y1"), self.P(sep2.tensors[0]), self._bnptr(bn1.name),
                                                     self._fin(bn1, Pn) if fuse_dw else None, b(p + "d2"), 0, 0, 0,
                                                     B, r, rw, f, 1, 0, 0, 0)))
            elif fuse_dw:
                prog.append((lib.crackfl_dw3x3_fwd_fin, (b(p + "y1"), self.P(sep2.tensors[0]), self._fin(bn1, Pn), b(p + "d2"),
                                                         B, r, rw, f)))
            else:
                prog.append((lib.crackfl_dw3x3_fwd, (b(p + "y1"), self.P(sep2.tensors[0]), self._bnptr(bn1.name),
                                                     b(p + "d2"), B, r, rw, f, 1)))
            sd = self._gemm(prog, b(p + "d2"), sep2.name + ".f", self.P(sep2.tensors[2]), 0, b(p + "y2"), B, r, rw,
                            self._zptr(bn2.name, "stats") if training else 0)
            bn(bn2, b(p + "y2"), Pn, sd, consumer_finalizes=fuse)
            prog.append((None, None, "join_aux"))
            last = e["k"] + 1 == len(self.enc)
            if fuse:
                prog.append((lib.crackfl_pool_add_fwd_fin, (b(p + "y2"), self._fin(bn2, Pn), b(p + "rs"), b(p + "out"),
                                                            0 if last else b(p + "oute"), b("a0_0") if last else 0,
                                                            b(p + "arg"), B, r, rw, f)))
            else:
                prog.append((lib.crackfl_pool_add_fwd, (b(p + "y2"), self._bnptr(bn2.name), b(p + "rs"), b(p + "out"),
                                                        0 if last else b(p + "oute"), b("a0_0") if last else 0,
                                                        b(p + "arg") if training else 0, B, r, rw, f)))
            xin, xe = p + "out", p + "oute"
        xprev = xin  # bottleneck tensor x3
        for d, (ct1, bn1, ct2, bn2, res) in zip(self.dec, decs):
            j, r, rw, cin, f = d["j"], d["r"], d["rw"], d["cin"], d["f"]
            p = f"d{j}."
            Pn = B * r * rw
            a0 = "a0_0" if j == 0 else p + "a0"
            lr_, lrw = (r, rw) if j == 0 else (r // 2, rw // 2)
            self._gemm(prog, b(xprev), res.name + ".f", self.P(res.tensors[1]), 0, b(p + "rl"), B, lr_, lrw, tag="aux")
            sd = self._gemm(prog, b(a0), ct1.name + ".f", self.P(ct1.tensors[1]), 0, b(p + "y1"), B, r, rw,
                            self._zptr(bn1.name, "stats") if training else 0)
            bn(bn1, b(p + "y1"), Pn, sd, consumer_finalizes=fuse)
            if self.conv_impl == "tc" and self.xf_dec and (f >= 64 or self.xf_dec >= 2) and r >= 16 and rw >= 8 and \
                    (fuse or not training) and self.gemm_stats:
                # K1: the second conv applies BN1 + ReLU to its input tiles itself (and writes the activation out for its
                # own weight gradient in training); no bn_act pass, no a1 read
                stats2 = self._zptr(bn2.name, "stats") if training else 0
                prog.append((lib.crackfl_conv_tc_xf, (b(p + "y1"), self.op_ptr(ct2.name + ".f"), self.P(ct2.tensors[1]), 0,
                                                      b(p + "y2"), stats2, B, r, rw, f, f, 2, self._bnptr(bn1.name),
                                                      self._fin(bn1, Pn) if training else None,
                                                      b(p + "a1") if training else 0, 0)))
                sd = training
            else:
                if fuse:
                    prog.append((lib.crackfl_bn_act_fin, (b(p + "y1"), self._fin(bn1, Pn), b(p + "a1"), 0, B, r, rw, f, 1)))
                else:
                    prog.append((lib.crackfl_bn_act, (b(p + "y1"), self._bnptr(bn1.name), b(p + "a1"), 0, B, r, rw, f, 1)))
                sd = self._gemm(prog, b(p + "a1"), ct2.name + ".f", self.P(ct2.tensors[1]), 0, b(p + "y2"), B, r, rw,
                                self._zptr(bn2.name, "stats") if training else 0)
            bn(bn2, b(p + "y2"), Pn, sd, consumer_finalizes=fuse)
            prog.append((None, None, "join_aux"))
            last = j + 1 == len(self.dec)
            if fuse:
                prog.append((lib.crackfl_dec_combine_fwd_fin, (b(p + "y2"), self._fin(bn2, Pn), b(p + "rl"), b(p + "s"),
                                                               0 if last else b(f"d{j + 1}.a0"), B, r, rw, f,
                                                               0 if j == 0 else 1)))
            else:
                prog.append((lib.crackfl_dec_combine_fwd, (b(p + "y2"), self._bnptr(bn2.name), b(p + "rl"), b(p + "s"),
                                                           0 if last else b(f"d{j + 1}.a0"), B, r, rw, f,
                                                           0 if j == 0 else 1)))
            xprev = p + "s"
        dl = self.dec[-1]
        prog.append((lib.crackfl_head_fwd, (b(xprev), self.P(head.tensors[0]), self.P(head.tensors[1]),
                                            b("mask_u8") if with_mask else 0, self._acc_ptr() if with_mask else 0,
                                            b("prob") if prob else 0, B, dl["r"], dl["rw"], dl["f"], self.bce_clip)))
        return prog

    def _build_train(self):
        lib, B, b = self.lib, self.B, self.b
        prog = self._build_forward(training=True, with_mask=True, prob=False)
        self.fwd_prog = list(prog)  # training-mode forward incl. the loss head (cross-checks, finite differences)
        self.n_fwd = len(prog)
        (conv0, bn0), encs, decs, head = self._layer_iter()
        ts = self.spec.tensors
        dl = self.dec[-1]
        last = f"d{dl['j']}."
        if not self.use_dice and os.environ.get("CRACKFL_HEAD_FUSED", "1") != "0":
            # plain BCE: the loss head's forward metrics and its backward in one pass (head_fwd is not launched)
            prog.pop()
            self.n_fwd = len(prog)
            prog.append((lib.crackfl_head_fwd_bwd, (b(last + "s"), self.P(head.tensors[0]), self.P(head.tensors[1]), b("mask_u8"),
                                                    self._acc_ptr(), b(last + "ds"), self.G(head.tensors[0]),
                                                    self.G(head.tensors[1]), B, dl["r"], dl["rw"], dl["f"], 1.0, self.bce_clip)))
        else:
            prog.append((lib.crackfl_head_bwd, (b(last + "s"), self.P(head.tensors[0]), self.P(head.tensors[1]), b("mask_u8"),
                                                self._acc_ptr(), b(last + "ds"), self.G(head.tensors[0]), self.G(head.tensors[1]),
                                                B, dl["r"], dl["rw"], dl["f"], self.use_dice, self.dice_smooth, 1.0,
                                                self.bce_clip)))
        # ---- decoder backward
        d2s_fused = {}
        for d, (ct1, bn1, ct2, bn2, res) in reversed(list(zip(self.dec, decs))):
            j, r, rw, cin, f = d["j"], d["r"], d["rw"], d["cin"], d["f"]
            p = f"d{j}."
            lr_, lrw = (r, rw) if j == 0 else (r // 2, rw // 2)
            src_res = self.enc_out_name() if j == 0 else f"d{j - 1}.s"
            fused_in = d2s_fused.get(j, False)  # ds, its BN2 reductions and drl came out of the previous block's tail launch
            if j > 0:
                if not fused_in:
                    prog.append((lib.crackfl_down2sum, (b(p + "ds"), 0, 0, b(p + "drl"), B, lr_, lrw, f)))
                drl = p + "drl"
            else:
                drl = p + "ds"
            # residual 1x1 conv (at low resolution): input gradient on the auxiliary stream, joined before the block tail
            self._gemm(prog, b(drl), res.name + ".d", 0, 0, b(p + "T"), B, lr_, lrw, tag="aux")
            # BN2 (no ReLU after it)
            if not fused_in:
                prog.append((lib.crackfl_bn_bwd_reduce, (b(p + "ds"), 0, b(p + "y2"), self._bnptr(bn2.name),
                                                         self._zptr(bn2.name, "sums"), B, r, rw, f, 0)))
            xf_b = self.conv_impl == "tc" and self.xf_bwd and (f >= 64 or self.xf_bwd >= 2) and r >= 16 and rw >= 8
            if xf_b:
                # K2: dy2 = BN-backward(ds, y2) is formed inside the input-gradient conv (and written out once for the
                # weight gradient); the launch also writes dgamma / dbeta
                prog.append((lib.crackfl_conv_tc_bnbwd, (b(p + "ds"), b(p + "y2"), self.op_ptr(ct2.name + ".d"), b(p + "da1"),
                                                         B, r, rw, f, f, self._bnptr(bn2.name), self._zptr(bn2.name, "sums"), 0,
                                                         b(p + "dy2"), self.G(bn2.tensors[0]), self.G(bn2.tensors[1]))))
            else:
                prog.append((lib.crackfl_bn_bwd_apply, (b(p + "ds"), 0, b(p + "y2"), self._bnptr(bn2.name),
                                                        self._zptr(bn2.name, "sums"), b(p + "dy2"), self.G(bn2.tensors[0]),
                                                        self.G(bn2.tensors[1]), B, r, rw, f, 0)))
                self._gemm(prog, b(p + "dy2"), ct2.name + ".d", 0, 0, b(p + "da1"), B, r, rw)
            oi = ts[ct2.tensors[0]]
            # Keras (kh,kw,out,in) kernel: with the operands swapped (A = dY, B = X) the tile columns run along "in", the
            # contiguous axis of the arena, so the split reduction writes coalesced; the tap flip then cancels out
            self._wgrad(prog, b(p + "dy2"), b(p + "a1"), self.G(ct2.tensors[0]), B, r, rw, f, f, 9, f * f, f, 1)
            # BN1 + ReLU
            prog.append((lib.crackfl_bn_bwd_reduce, (b(p + "da1"), 0, b(p + "y1"), self._bnptr(bn1.name),
                                                     self._zptr(bn1.name, "sums"), B, r, rw, f, 1)))
            a0 = "a0_0" if j == 0 else p + "a0"
            if xf_b:
                prog.append((lib.crackfl_conv_tc_bnbwd, (b(p + "da1"), b(p + "y1"), self.op_ptr(ct1.name + ".d"), b(p + "da0"),
                                                         B, r, rw, f, cin, self._bnptr(bn1.name), self._zptr(bn1.name, "sums"), 1,
                                                         b(p + "dy1"), self.G(bn1.tensors[0]), self.G(bn1.tensors[1]))))
            else:
                prog.append((lib.crackfl_bn_bwd_apply, (b(p + "da1"), 0, b(p + "y1"), self._bnptr(bn1.name),
                                                        self._zptr(bn1.name, "sums"), b(p + "dy1"), self.G(bn1.tensors[0]),
                                                        self.G(bn1.tensors[1]), B, r, rw, f, 1)))
                self._gemm(prog, b(p + "dy1"), ct1.name + ".d", 0, 0, b(p + "da0"), B, r, rw)
            self._wgrad(prog, b(p + "dy1"), b(a0), self.G(ct1.tensors[0]), B, r, rw, f, cin, 9, f * cin, cin, 1)
            self._wgrad(prog, b(src_res), b(drl), self.G(res.tensors[0]), B, lr_, lrw, cin, f, 1, 0, f, 1)
            prog.append((lib.crackfl_colsum, (b(drl), self.G(res.tensors[1]), B * lr_ * lrw, f), os.environ.get("CRACKFL_COLSUM_STREAM", "side3")))
            prog.append((None, None, "join_aux"))
            self._wgrad_flush(prog)
            if j > 0 and self.d2s_bn and lr_ % 2 == 0 and lrw % 2 == 0 and 256 % (cin // 8) == 0:
                # one launch: ds of the block below, the BN2-backward reductions over it and its own down2sum (drl)
                pbn2 = decs[j - 1][3]
                prog.append((lib.crackfl_down2sum_bn, (b(p + "da0"), b(f"d{j - 1}.s"), b(p + "T"), b(f"d{j - 1}.ds"),
                                                       b(f"d{j - 1}.drl") if j - 1 > 0 else 0, b(f"d{j - 1}.y2"),
                                                       self._bnptr(pbn2.name), self._zptr(pbn2.name, "sums"), B, lr_, lrw, cin)))
                d2s_fused[j - 1] = True
            elif j > 0:
                prog.append((lib.crackfl_down2sum, (b(p + "da0"), b(f"d{j - 1}.s"), b(p + "T"), b(f"d{j - 1}.ds"),
                                                    B, lr_, lrw, cin)))
            else:
                prog.append((lib.crackfl_relu_bwd_add, (b(p + "da0"), b(src_res), b(p + "T"), b("dx_bott"),
                                                        B * r * rw * cin)))
        # ---- the decoder's and the head's gradients are complete once the side streams drain: their Adam update runs
        # now, next to the encoder's backward pass, and the launch at the end of the step only has the encoder's parameters
        # left (the bf16 operand copies, not the arena, feed the rest of this step's GEMMs; the depthwise kernels and the
        # BN tables the encoder's backward reads are outside this range)
        adam_split = 0
        if self.adam_early and decs:
            first = min(ts[t].offset for t in decs[0][0].tensors)
            adam_split = (first + 3) // 4 * 4
            cs = os.environ.get("CRACKFL_COLSUM_STREAM", "side3")
            if cs in ("side2", "side3"):
                prog.append((None, None, ("rec", cs, "dec_colsum")))
                prog.append((None, None, ("wait", "side", "dec_colsum")))
            n_early = self.spec.arena_numel - adam_split
            prog.append((lib.crackfl_adam_part, (self.arena.data_ptr() + 4 * adam_split, self.grads.data_ptr() + 4 * adam_split,
                                                 self.m.data_ptr() + 4 * adam_split, self.v.data_ptr() + 4 * adam_split,
                                                 self.step.data_ptr(), n_early, self.lr, self.b1, self.b2, self.eps), "side"))
        # ---- encoder backward
        dx_next = "dx_bott"
        for e, (sep1, bn1, sep2, bn2, res) in reversed(list(zip(self.enc, encs))):
            k, r, rw, cin, f = e["k"], e["r"], e["rw"], e["cin"], e["f"]
            p = f"e{k}."
            xin = "x0" if k == 0 else f"e{k - 1}.out"
            xe = "x0e" if k == 0 else f"e{k - 1}.oute"
            r2, rw2 = r // 2, rw // 2
            # residual 1x1 s2 conv
            self._gemm(prog, b(dx_next), res.name + ".d", 0, 0, b(p + "Te"), B, r2, rw2, tag="aux")
            self._wgrad(prog, b(xe), b(dx_next), self.G(res.tensors[0]), B, r2, rw2, cin, f, 1, 0, f, 1)
            prog.append((lib.crackfl_colsum, (b(dx_next), self.G(res.tensors[1]), B * r2 * rw2, f), os.environ.get("CRACKFL_COLSUM_STREAM", "side3")))
            # max-pool + BN2
            # max-pool backward once into dd2 (reused below as the pointwise dgrad output), then DIRECT-mode BN backward
            if self.pool_bwd_fused:
                prog.append((lib.crackfl_pool_bwd_bn, (b(dx_next), b(p + "arg"), b(p + "y2"), self._bnptr(bn2.name), b(p + "dd2"),
                                                       self._zptr(bn2.name, "sums"), B, r, rw, f)))
            else:
                prog.append((lib.crackfl_pool_bwd, (b(dx_next), b(p + "arg"), b(p + "dd2"), B, r, rw, f)))
                prog.append((lib.crackfl_bn_bwd_reduce, (b(p + "dd2"), 0, b(p + "y2"), self._bnptr(bn2.name),
                                                         self._zptr(bn2.name, "sums"), B, r, rw, f, 0)))
            prog.append((lib.crackfl_bn_bwd_apply, (b(p + "dd2"), 0, b(p + "y2"), self._bnptr(bn2.name),
                                                    self._zptr(bn2.name, "sums"), b(p + "dy2"), self.G(bn2.tensors[0]),
                                                    self.G(bn2.tensors[1]), B, r, rw, f, 0)))
            # pointwise 2
            self._gemm(prog, b(p + "dy2"), sep2.name + ".d", 0, 0, b(p + "dd2"), B, r, rw)
            self._wgrad(prog, b(p + "d2"), b(p + "dy2"), self.G(sep2.tensors[1]), B, r, rw, f, f, 1, 0, f, 1)
            # depthwise 2 (input = relu(bn1(y1)))
            folded = self._dw_bwd(prog, b(p + "dd2"), b(p + "y1"), self.P(sep2.tensors[0]), self._bnptr(bn1.name), 0,
                                  b(p + "dz1"), self.G(sep2.tensors[0]), B, r, rw, f, 1, 0,
                                  red_bn=self._bnptr(bn1.name), red_sums=self._zptr(bn1.name, "sums"))
            if not folded:
                prog.append((lib.crackfl_bn_bwd_reduce, (b(p + "dz1"), 0, b(p + "y1"), self._bnptr(bn1.name),
                                                         self._zptr(bn1.name, "sums"), B, r, rw, f, 0)))
            prog.append((lib.crackfl_bn_bwd_apply, (b(p + "dz1"), 0, b(p + "y1"), self._bnptr(bn1.name),
                                                    self._zptr(bn1.name, "sums"), b(p + "dy1"), self.G(bn1.tensors[0]),
                                                    self.G(bn1.tensors[1]), B, r, rw, f, 0)))
            # pointwise 1
            self._gemm(prog, b(p + "dy1"), sep1.name + ".d", 0, 0, b(p + "dd1"), B, r, rw)
            self._wgrad(prog, b(p + "d1"), b(p + "dy1"), self.G(sep1.tensors[1]), B, r, rw, cin, f, 1, 0, f, 1)
            # depthwise 1 (input = relu(x_k)); adds the residual-path gradient on the even pixels
            out = "dz0" if k == 0 else p + "dx"
            if k > 0:
                self._wgrad_flush(prog)  # (the first block's batch also takes the entry conv, flushed before Adam)
            prog.append((None, None, "join_aux"))
            folded0 = self._dw_bwd(prog, b(p + "dd1"), b(xin), self.P(sep1.tensors[0]), 0, b(p + "Te"), b(out),
                                   self.G(sep1.tensors[0]), B, r, rw, cin, 0, 2 if k == 0 else 1,
                                   red_bn=self._bnptr(bn0.name) if k == 0 else 0,
                                   red_sums=self._zptr(bn0.name, "sums") if k == 0 else 0)
            dx_next = out
        # ---- entry
        r, rw, c0 = self.H // 2, self.W // 2, self.spec.entry
        if not folded0:  # (folded0 is from the k == 0 block, the last one of the reversed loop)
            prog.append((lib.crackfl_bn_bwd_reduce, (b("dz0"), 0, b("y0"), self._bnptr(bn0.name), self._zptr(bn0.name, "sums"),
                                                     B, r, rw, c0, 0)))
        prog.append((lib.crackfl_bn_bwd_apply, (b("dz0"), 0, b("y0"), self._bnptr(bn0.name), self._zptr(bn0.name, "sums"),
                                                b("dy0"), self.G(bn0.tensors[0]), self.G(bn0.tensors[1]), B, r, rw, c0, 0)))
        if self.entry_tc:
            self._wgrad(prog, b("x0col"), b("dy0"), self.G(conv0.tensors[0]), B, r, rw, 32, c0, 1, 0, c0, 1, cin_valid=27)
        else:
            prog.append((lib.crackfl_entry_conv_wgrad, (b("x_u8"), b("dy0"), self.G(conv0.tensors[0]), B, self.H, self.W, c0),
                         "side"))
        self._wgrad_flush(prog)
        prog.append((None, None, "join"))
        # ---- optimiser (also zeroes the gradient arena)
        # Adam, its step tick and the metric accumulation in one launch (counter slot acc_off + 8 of the zeroed region)
        prog.append((lib.crackfl_adam_step_tail, (self.arena.data_ptr(), self.grads.data_ptr(), self.m.data_ptr(),
                                                  self.v.data_ptr(), self.step.data_ptr(),
                                                  adam_split if adam_split else self.spec.arena_numel, self.lr,
                                                  self.b1, self.b2, self.eps, self._acc_ptr() + 4 * 8,
                                                  self.metrics_total.data_ptr(), self._acc_ptr(), 6,
                                                  self.last_acc.data_ptr(), self.zero_region.data_ptr(),
                                                  self.zero_region.numel())))
        return self._wgrad_resolve(prog)

    def enc_out_name(self) -> str:
        return f"e{len(self.enc) - 1}.out"

    # ------------------------------------------------------------------ execution
    def _run(self, prog):
        """Execute a schedule.  Entries tagged "side" (weight / bias gradients: nothing downstream but Adam reads
        them) are forked onto a second stream after the main-stream work issued so far, so they overlap with the
        dgrad chain; "join" makes the main stream wait for all of them.  Entries tagged "aux" (the small residual-path
        GEMMs) fork onto a third stream the same way and are joined by the next "join_aux" marker, placed before their
        consumer.  Works eagerly and under graph capture."""
        main = torch.cuda.current_stream(self.dev)
        st = main.cuda_stream
        streams = {"side": self._side_stream, "aux": self._aux_stream, "side2": self._side2_stream,
                   "side3": self._side3_stream}
        forked = {"side": False, "aux": False, "side2": False, "side3": False}
        events = {}

        def join(name):
            if forked[name]:
                ev = torch.cuda.Event()
                ev.record(streams[name])
                main.wait_event(ev)
                forked[name] = False

        for entry in prog:
            fn, args = entry[0], entry[1]
            tag = entry[2] if len(entry) > 2 else None
            if isinstance(tag, tuple):  # ("rec" | "wait", stream, key): explicit event edges between the extra streams
                if self.side_streams:
                    s_ = main if tag[1] == "main" else streams[tag[1]]
                    if tag[0] == "rec" and s_ is not main and not forked.get(tag[1], False):
                        continue  # nothing was launched on that stream since the last join: there is nothing to wait for
                    if tag[0] == "wait" and tag[2] not in events:
                        continue
                    if tag[0] == "rec":
                        events[tag[2]] = torch.cuda.Event()
                        events[tag[2]].record(s_)
                    else:
                        s_.wait_event(events[tag[2]])
                continue
            if tag == "join":
                join("side")
                join("side2")
                join("side3")
                continue
            if tag == "join_aux":
                join("aux")
                continue
            if tag in streams and self.side_streams:
                if tag != "side2":  # side2 work is ordered by explicit events only
                    ev = torch.cuda.Event()
                    ev.record(main)
                    streams[tag].wait_event(ev)
                rc = fn(*args, streams[tag].cuda_stream)
                forked[tag] = True
            else:
                rc = fn(*args, st)
            if rc != 0:
                raise L.KernelError(f"{fn.__name__} failed with cudaError {rc}")
        join("side")
        join("side2")
        join("side3")
        join("aux")

    def _step_body(self, prog, accumulate: bool):
        """The per-step scratch region (BN sums, loss accumulators, the Adam tail counter) must be zero when a step
        starts.  A train step leaves it that way itself (last block of the Adam launch), so the hot path has no fill
        kernel; eval / predict steps clear it before they run (``_exec`` clears it again before the next train step)."""
        if prog is not self.train_prog:
            self.zero_region.zero_()
        self._run(prog)
        if accumulate and prog is not self.train_prog:  # the train schedule accumulates inside its Adam launch
            st = torch.cuda.current_stream(self.dev).cuda_stream
            L.check(self.lib.crackfl_accum_small(self.metrics_total.data_ptr(), self._acc_ptr(), 6, st))

    def _exec(self, key: str, prog, accumulate: bool):
        with torch.cuda.device(self.dev):
            if key == "train":
                if self._region_dirty:
                    self.zero_region.zero_()
                    self._region_dirty = False
            else:
                self._region_dirty = True
            if not self.cuda_graph:
                self._step_body(prog, accumulate)
                return
            g = self._graphs.get(key)
            if g is None:
                # one eager run first (lazy module loading / cudaFuncSetAttribute must not happen during capture)
                saved = None
                if key == "train":
                    saved = (self.arena.clone(), self.m.clone(), self.v.clone(), self.step.clone(),
                             self.metrics_total.clone(), self.grads.clone())
                else:
                    saved = (self.metrics_total.clone(),)
                torch.cuda.synchronize(self.dev)
                self._step_body(prog, accumulate)
                torch.cuda.synchronize(self.dev)
                if key == "train":
                    self.arena.copy_(saved[0]); self.m.copy_(saved[1]); self.v.copy_(saved[2]); self.step.copy_(saved[3])
                    self.metrics_total.copy_(saved[4]); self.grads.copy_(saved[5])
                else:
                    self.metrics_total.copy_(saved[0])
                g = torch.cuda.CUDAGraph()
                s = torch.cuda.Stream(self.dev, priority=self._main_prio)
                s.wait_stream(torch.cuda.current_stream(self.dev))
                with torch.cuda.stream(s):
                    with torch.cuda.graph(g, stream=s):
                        self._step_body(prog, accumulate)
                torch.cuda.current_stream(self.dev).wait_stream(s)
                self._graphs[key] = g
                # the capture itself did not execute anything; fall through to replay
            g.replay()

    def profile_step(self, repeats: int = 3):
        """In-situ per-kernel timing: the train schedule run eagerly on ONE stream with a CUDA event after every
        launch (warm L2, real producer/consumer adjacency - unlike ncu's serialised cold-cache replay).
        Returns [(kernel name, ms)] averaged over ``repeats`` steps."""
        names = [e[0].__name__ for e in self.train_prog if e[0] is not None]
        acc = [0.0] * len(names)
        st = torch.cuda.current_stream(self.dev)
        self._region_dirty = True
        for rep in range(repeats + 1):
            self.zero_region.zero_()
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
            evs[0].record(st)
            i = 0
            for e in self.train_prog:
                if e[0] is None:
                    continue
                rc = e[0](*e[1], st.cuda_stream)
                if rc:
                    raise L.KernelError(f"{e[0].__name__}: {rc}")
                i += 1
                evs[i].record(st)
            st.synchronize()
            if rep > 0:
                for k in range(len(names)):
                    acc[k] += evs[k].elapsed_time(evs[k + 1]) / repeats
        return list(zip(names, acc))

    # ------------------------------------------------------------------ public API
    def load_weights(self, weights: Sequence[np.ndarray]):
        flat = torch.from_numpy(wio.flatten(weights, self.spec))
        self.arena.copy_(flat.to(self.dev, non_blocking=False))

    def load_flat(self, flat: torch.Tensor):
        self.arena.copy_(flat)

    def get_weights(self) -> List[np.ndarray]:
        return wio.unflatten(self.arena.detach().cpu().numpy(), self.spec)

    def reset_optimizer(self):
        """Fresh ``compile()``: Adam moments and step count to zero (client_fit_model.py:155-157)."""
        self.m.zero_(); self.v.zero_(); self.step.zero_(); self.grads.zero_()

    def reset_metrics(self):
        self.metrics_total.zero_()

    def read_metrics(self) -> Dict[str, float]:
        t = self.metrics_total.detach().cpu().numpy()
        n = max(float(t[5]), 1.0)
        out = {"loss": float(t[0]) / n, "accuracy": float(t[1]) / n}
        if self.use_dice:
            # per-batch Dice averaged approximately via the accumulated sums
            out["dice"] = float((2 * t[2] + self.dice_smooth) / (t[3] + t[4] + self.dice_smooth))
        return out

    def set_batch(self, x_u8, y_u8=None):
        """x: (B,H,W,3) uint8 numpy/torch (host or device); y: (B,H,W[,1]) uint8."""
        xt = torch.as_tensor(x_u8)
        self.x_u8.copy_(xt.view(self.x_u8.shape) if xt.dtype == torch.uint8 else xt.to(torch.uint8), non_blocking=True)
        if y_u8 is not None:
            yt = torch.as_tensor(y_u8)
            self.mask_u8.copy_(yt.reshape(self.mask_u8.shape), non_blocking=True)

    def train_step(self, x_u8=None, y_u8=None):
        if x_u8 is not None:
            self.set_batch(x_u8, y_u8)
        self._exec("train", self.train_prog, True)

    # -- input prefetch: the H2D copy of batch i+1 runs on a copy stream while batch i trains -----------------
    def prefetch(self, x_u8, y_u8):
        """Start the host->device copy of the NEXT batch (pinned host tensors) on the copy stream."""
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream(self.dev)
            self._stage = [(torch.empty_like(self.x_u8), torch.empty_like(self.mask_u8)) for _ in range(2)]
            self._stage_ev = [torch.cuda.Event(), torch.cuda.Event()]
            self._stage_free = [torch.cuda.Event(), torch.cuda.Event()]
            self._stage_i = 0
            self._stage_pending = None
        i = self._stage_i
        sx, sy = self._stage[i]
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._stage_free[i])  # the step that consumed this slot has read it
            sx.copy_(torch.as_tensor(x_u8).view(sx.shape), non_blocking=True)
            sy.copy_(torch.as_tensor(y_u8).reshape(sy.shape), non_blocking=True)
            self._stage_ev[i].record(self._copy_stream)
        self._stage_pending = i
        self._stage_i ^= 1

    def train_step_prefetched(self):
        """Train on the batch handed to the last ``prefetch`` call (device-to-device move of 4 MB, then the graph)."""
        i = self._stage_pending
        if i is None:
            raise RuntimeError("prefetch() a batch first")
        main = torch.cuda.current_stream(self.dev)
        main.wait_event(self._stage_ev[i])
        self.x_u8.copy_(self._stage[i][0], non_blocking=True)
        self.mask_u8.copy_(self._stage[i][1], non_blocking=True)
        self._stage_free[i].record(main)
        self._stage_pending = None
        self._exec("train", self.train_prog, True)

    def eval_step(self, x_u8=None, y_u8=None):
        if x_u8 is not None:
            self.set_batch(x_u8, y_u8)
        self._exec("eval", self.eval_prog, True)

    def predict(self, x_u8=None) -> torch.Tensor:
        """Returns probabilities (B,H,W) fp32 on the device (valid until the next call)."""
        if x_u8 is not None:
            self.set_batch(x_u8)
        self._exec("pred", self.pred_prog, False)
        return self.prob

    def step_loss(self) -> torch.Tensor:
        """Device scalar: mean BCE of the most recent step (for the end-to-end D2H read)."""
        a = self.last_acc[:6]
        return a[0] / a[5].clamp_min(1.0)


    def step_loss_host(self) -> float:
        """Mean BCE of the most recent step read back to the host: ONE 24-byte D2H copy of the step's accumulators into
        pinned memory + a stream synchronise (no device-side arithmetic kernels), the division is done on the host."""
        if getattr(self, "_acc_host", None) is None:
            self._acc_host = torch.empty(6, dtype=torch.float32, pin_memory=True)
        self._acc_host.copy_(self.last_acc[:6], non_blocking=True)
        torch.cuda.current_stream(self.dev).synchronize()
        a = self._acc_host
        return float(a[0]) / max(float(a[5]), 1.0)


    def enqueue_loss_read(self) -> int:
        """Asynchronous form of ``step_loss_host``: enqueue the 24-byte D2H copy of the step that was just launched (stream
        ordered before the next step clears the accumulators) and return a ticket for ``wait_loss``.  Lets the caller
        launch step i+1 before it blocks on the loss of step i, so the GPU never idles on the host's launch latency."""
        if getattr(self, "_lr_host", None) is None:
            self._lr_host = [torch.empty(6, dtype=torch.float32, pin_memory=True) for _ in range(2)]
            self._lr_ev = [torch.cuda.Event(), torch.cuda.Event()]
            self._lr_i = 0
        i = self._lr_i
        self._lr_i ^= 1
        self._lr_host[i].copy_(self.last_acc[:6], non_blocking=True)
        self._lr_ev[i].record(torch.cuda.current_stream(self.dev))
        return i

    def wait_loss(self, ticket: int) -> float:
        self._lr_ev[ticket].synchronize()
        a = self._lr_host[ticket]
        return float(a[0]) / max(float(a[5]), 1.0)


class B200Backend:
    """``trainer.LocalTrainer`` back-end on the engine (one engine per (batch, size) shape)."""

    name = "b200"

    def __init__(self, cfg, spec: UNetSpec, device: Optional[int] = None, arena: Optional[torch.Tensor] = None):
        t = cfg.train
        self.cfg, self.spec = cfg, spec
        self.engine = UNetEngine(spec, t.batch_size, t.image_size[0], t.image_size[1], device=device, lr=t.lr,
                                 beta1=t.beta1, beta2=t.beta2, eps=t.adam_eps, loss=t.loss, cuda_graph=t.cuda_graph,
                                 bce_clip=t.bce_clip, arena=arena)
        B, H, W = t.batch_size, t.image_size[0], t.image_size[1]
        self._hx = torch.empty((B, H, W, 3), dtype=torch.uint8).pin_memory()
        self._hy = torch.empty((B, H, W), dtype=torch.uint8).pin_memory()

    def begin_round(self, weights):
        if weights is not None:  # None: the round's model is already in the arena (on-box broadcast)
            self.engine.load_weights(weights)
        self.engine.reset_optimizer()
        self.engine.reset_metrics()

    def reset_metrics(self):
        self.engine.reset_metrics()

    def _stage(self, x_u8, y_u8):
        if torch.is_tensor(x_u8) and x_u8.is_cuda:  # device-resident data (synthetic generator K15): no staging
            return x_u8, y_u8
        torch.cuda.current_stream(self.engine.dev).synchronize()  # previous H2D from the pinned buffers is done
        np.copyto(self._hx.numpy(), np.asarray(x_u8).reshape(self._hx.shape), casting="unsafe")
        if y_u8 is not None:
            np.copyto(self._hy.numpy(), np.asarray(y_u8).reshape(self._hy.shape), casting="unsafe")
        return self._hx, (self._hy if y_u8 is not None else None)

    def train_batch(self, x_u8, y_u8):
        """Host batches go through two pinned buffers in turn and the engine's copy stream, so preparing batch i+1 on the
        host (and its H2D copy) overlaps training on batch i; only the copy that last READ a pinned buffer is waited for,
        never the training stream."""
        if torch.is_tensor(x_u8) and x_u8.is_cuda:
            self.engine.train_step(x_u8, y_u8)
            return
        eng = self.engine
        if not hasattr(self, "_pin"):
            self._pin = [(torch.empty_like(self._hx).pin_memory(), torch.empty_like(self._hy).pin_memory()) for _ in range(2)]
            self._pin_ev = [None, None]
            self._pi = 0
        i = self._pi
        self._pi ^= 1
        if self._pin_ev[i] is not None:
            self._pin_ev[i].synchronize()  # the H2D copy that read this pinned buffer two batches ago is done
        hx, hy = self._pin[i]
        # plain memcpy through NumPy views of the pinned buffers: torch's CPU copy_ splits a 3 MB copy over its OpenMP
        # pool, and with decode threads busy on the same cores the slowest slice set the pace (measured: 9.5 ms per
        # batch in train_batch against 0.2 ms for the memcpy)
        np.copyto(hx.numpy(), np.asarray(x_u8).reshape(hx.shape), casting="unsafe")
        np.copyto(hy.numpy(), np.asarray(y_u8).reshape(hy.shape), casting="unsafe")
        eng.prefetch(hx, hy)
        self._pin_ev[i] = eng._stage_ev[eng._stage_pending]
        eng.train_step_prefetched()

    def eval_batch(self, x_u8, y_u8):
        self.engine.eval_step(*self._stage(x_u8, y_u8))

    def read_metrics(self):
        return self.engine.read_metrics()

    def finish(self):
        torch.cuda.current_stream(self.engine.dev).synchronize()

    def predict_batch(self, x_u8) -> np.ndarray:
        hx, _ = self._stage(x_u8, None)
        return self.engine.predict(hx).detach().cpu().numpy()[..., None]

    def get_weights(self):
        return self.engine.get_weights()
