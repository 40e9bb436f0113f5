"""R-train stand-in: the reference's U-Net graph (client_fit_model.py:92-150) in STOCK PyTorch
eager mode - cuDNN/cuBLAS kernels, bf16 autocast, channels_last, torch.optim.Adam(fused) - because TensorFlow/Keras
cannot be installed on this image.  It is a stand-in for "what a library-kernel build achieves on this box",
NOT the reference's own build, and is always labelled so.  None of crackfl's kernels are on this path.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class SepConv(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.dw = nn.Conv2d(cin, cin, 3, padding=1, groups=cin, bias=False)
        self.pw = nn.Conv2d(cin, cout, 1)

    def forward(self, x):
        return self.pw(self.dw(x))


class StandinUNet(nn.Module):
    def __init__(self, spec):
        super().__init__()
        bn = lambda c: nn.BatchNorm2d(c, eps=spec.bn_eps, momentum=1 - spec.bn_momentum)
        self.entry = nn.Conv2d(spec.in_channels, spec.entry, 3, stride=2)
        self.entry_bn = bn(spec.entry)
        self.down = nn.ModuleList()
        prev = spec.entry
        for f in spec.down:
            self.down.append(nn.ModuleList([SepConv(prev, f), bn(f), SepConv(f, f), bn(f), nn.Conv2d(prev, f, 1, stride=2)]))
            prev = f
        self.up = nn.ModuleList()
        for f in spec.up:
            self.up.append(nn.ModuleList([nn.ConvTranspose2d(prev, f, 3, padding=1), bn(f),
                                          nn.ConvTranspose2d(f, f, 3, padding=1), bn(f), nn.Conv2d(prev, f, 1)]))
            prev = f
        self.head = nn.Conv2d(prev, 1, 1)

    def forward(self, x):
        x = F.relu(self.entry_bn(self.entry(F.pad(x, (0, 1, 0, 1)))))
        prev = x
        for s1, b1, s2, b2, res in self.down:
            x = b1(s1(F.relu(x)))
            x = b2(s2(F.relu(x)))
            x = F.max_pool2d(F.pad(x, (0, 1, 0, 1), value=float("-inf")), 3, 2)
            x = x + res(prev)
            prev = x
        for c1, b1, c2, b2, res in self.up:
            x = b1(c1(F.relu(x)))
            x = b2(c2(F.relu(x)))
            x = F.interpolate(x, scale_factor=2, mode="nearest")
            x = x + res(F.interpolate(prev, scale_factor=2, mode="nearest"))
            prev = x
        return self.head(x)


class StandinTrainer:
    """mode: "eager" (Python-driven launches), "graph" (the whole step - zero_grad, forward, loss, backward, fused Adam -
    captured once in a CUDA graph and replayed: no launch overhead, the fair comparison for a graph-replayed engine),
    "compile" (torch.compile of the model, Inductor/Triton fusions, still Python-driven optimiser).
    dtype: "bf16" (autocast) or "fp32" (what the reference trains in, client_fit_model.py:157; cuDNN may use TF32)."""

    def __init__(self, spec, device, lr=1e-3, autocast=True, mode="eager", dtype="bf16", batch_shape=None):
        self.dev = device
        self.mode = mode
        self.model = StandinUNet(spec).to(device).to(memory_format=torch.channels_last)
        self.autocast = autocast and dtype == "bf16"
        self.opt = torch.optim.Adam(self.model.parameters(), lr=lr, eps=1e-7, fused=True, capturable=(mode == "graph"))
        self.fwd = torch.compile(self.model) if mode == "compile" else self.model
        self._graph = None
        self._sx = self._sy = self._sloss = None

    def _step(self, x_u8, y_u8):
        x = x_u8.permute(0, 3, 1, 2).float().div_(255.0).contiguous(memory_format=torch.channels_last)
        y = y_u8.unsqueeze(1).float()
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.autocast):
            logits = self.fwd(x)
        loss = F.binary_cross_entropy_with_logits(logits.float(), y)
        loss.backward()
        self.opt.step()
        return loss

    def _capture(self, x_u8, y_u8):
        self._sx, self._sy = x_u8.clone(), y_u8.clone()
        s = torch.cuda.Stream(self.dev)
        s.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(s):
            for _ in range(3):  # warm-up on a side stream (cuDNN autotune, lazy state) before capture
                self._step(self._sx, self._sy)
        torch.cuda.current_stream(self.dev).wait_stream(s)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._sloss = self._step(self._sx, self._sy)

    def _graph_step(self, x_u8, y_u8):
        if self._graph is None:
            self._capture(x_u8, y_u8)
        self._sx.copy_(x_u8, non_blocking=True)
        self._sy.copy_(y_u8, non_blocking=True)
        self._graph.replay()
        return self._sloss

    def train_step_device(self, x_u8_dev, y_u8_dev):
        if self.mode == "graph":
            return self._graph_step(x_u8_dev, y_u8_dev)
        return self._step(x_u8_dev, y_u8_dev)

    def train_step_host(self, x_u8_pinned, y_u8_pinned):
        xd, yd = x_u8_pinned.to(self.dev, non_blocking=True), y_u8_pinned.to(self.dev, non_blocking=True)
        loss = self._graph_step(xd, yd) if self.mode == "graph" else self._step(xd, yd)
        return float(loss.item())
