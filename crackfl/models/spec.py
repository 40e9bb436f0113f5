"""Canonical description of the crack-segmentation U-Net and its 112 weight tensors.

The graph is the Xception-style encoder/decoder built by the reference at
``client_fit_model.py:92-150`` (duplicated at ``test/Segmentation.py:102-159``):

    entry   Conv3x3 s2 (3->32) + BN + ReLU
    down_k  [ReLU, SepConv3x3, BN] x2, MaxPool3x3 s2, + Conv1x1 s2 residual   f = 64,128,256
    up_k    [ReLU, ConvT3x3, BN] x2, Up2x, + Conv1x1(Up2x(prev)) residual     f = 256,128,64,32
    head    Conv1x1 (->1) + sigmoid

Weights are kept in the *Keras layouts* (what ``model.get_weights()`` returns and what
the reference pickles onto the wire, ``fl_client.py:63``), in a fixed construction
order defined here.  Everything else in the framework (flat arena, FedAvg, pickle, h5,
native checkpoints) is a view over this table.

``widths`` generalises the model ("wide U-Net" in BASELINE.json config 4 is a
north-star extension, not something the reference has).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import numpy as np

# Reference defaults (client_fit_model.py:100,107,127)
ENTRY_FILTERS = 32
DOWN_FILTERS = (64, 128, 256)
UP_FILTERS = (256, 128, 64, 32)
BN_MOMENTUM = 0.99  # Keras default
BN_EPS = 1e-3  # Keras default


@dataclass(frozen=True)
class TensorSpec:
    """One entry of ``get_weights()``."""

    layer: str  # Keras layer name, e.g. "separable_conv2d_3"
    weight: str  # Keras weight name without the ":0", e.g. "depthwise_kernel"
    shape: Tuple[int, ...]
    trainable: bool
    kind: str  # conv_kernel | bias | dw_kernel | pw_kernel | convT_kernel | gamma | beta | mean | var
    offset: int = 0  # element offset in the flat arena

    @property
    def name(self) -> str:
        return f"{self.layer}/{self.weight}:0"

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


@dataclass(frozen=True)
class LayerSpec:
    name: str
    op: str  # conv | bn | sepconv | convT
    cin: int
    cout: int
    ksize: int = 1
    stride: int = 1
    tensors: Tuple[int, ...] = ()  # indices into UNetSpec.tensors


@dataclass
class UNetSpec:
    """Full model description; ``UNetSpec()`` is the reference model (2,058,145 params)."""

    in_channels: int = 3
    entry: int = ENTRY_FILTERS
    down: Sequence[int] = DOWN_FILTERS
    up: Sequence[int] = UP_FILTERS
    bn_momentum: float = BN_MOMENTUM
    bn_eps: float = BN_EPS
    tensors: List[TensorSpec] = field(default_factory=list, init=False)
    layers: List[LayerSpec] = field(default_factory=list, init=False)

    def __post_init__(self):
        self.down = tuple(self.down)
        self.up = tuple(self.up)
        if len(self.up) != len(self.down) + 1:
            raise ValueError("decoder must have one more block than the encoder (stride-2 entry conv)")
        self._build()

    # -- construction -----------------------------------------------------------------
    def _add(self, layer: str, op: str, cin: int, cout: int, ksize: int, stride: int, ts):
        idx = []
        for weight, shape, trainable, kind in ts:
            off = self.tensors[-1].offset + self.tensors[-1].numel if self.tensors else 0
            self.tensors.append(TensorSpec(layer, weight, tuple(shape), trainable, kind, off))
            idx.append(len(self.tensors) - 1)
        self.layers.append(LayerSpec(layer, op, cin, cout, ksize, stride, tuple(idx)))

    def _build(self):
        counters: Dict[str, int] = {}

        def nm(base: str) -> str:
            k = counters.get(base, 0)
            counters[base] = k + 1
            return base if k == 0 else f"{base}_{k}"

        def conv(cin, cout, k, s):
            self._add(nm("conv2d"), "conv", cin, cout, k, s,
                      [("kernel", (k, k, cin, cout), True, "conv_kernel"), ("bias", (cout,), True, "bias")])

        def bn(c):
            self._add(nm("batch_normalization"), "bn", c, c, 1, 1,
                      [("gamma", (c,), True, "gamma"), ("beta", (c,), True, "beta"),
                       ("moving_mean", (c,), False, "mean"), ("moving_variance", (c,), False, "var")])

        def sep(cin, cout):
            self._add(nm("separable_conv2d"), "sepconv", cin, cout, 3, 1,
                      [("depthwise_kernel", (3, 3, cin, 1), True, "dw_kernel"),
                       ("pointwise_kernel", (1, 1, cin, cout), True, "pw_kernel"),
                       ("bias", (cout,), True, "bias")])

        def convT(cin, cout):
            # Keras Conv2DTranspose stores (kh, kw, out, in)
            self._add(nm("conv2d_transpose"), "convT", cin, cout, 3, 1,
                      [("kernel", (3, 3, cout, cin), True, "convT_kernel"), ("bias", (cout,), True, "bias")])

        conv(self.in_channels, self.entry, 3, 2)
        bn(self.entry)
        prev = self.entry
        for f in self.down:
            sep(prev, f); bn(f)
            sep(f, f); bn(f)
            conv(prev, f, 1, 2)
            prev = f
        for f in self.up:
            convT(prev, f); bn(f)
            convT(f, f); bn(f)
            conv(prev, f, 1, 1)
            prev = f
        conv(prev, 1, 1, 1)

    # -- queries ----------------------------------------------------------------------
    @property
    def num_tensors(self) -> int:
        return len(self.tensors)

    @property
    def num_params(self) -> int:
        return sum(t.numel for t in self.tensors)

    @property
    def num_trainable(self) -> int:
        return sum(t.numel for t in self.tensors if t.trainable)

    @property
    def arena_numel(self) -> int:
        """Flat-arena length: parameters padded with zeros to a multiple of 1024 elements."""
        n = self.num_params
        return (n + 1023) // 1024 * 1024

    def index(self, name: str) -> int:
        for i, t in enumerate(self.tensors):
            if t.name == name or f"{t.layer}/{t.weight}" == name:
                return i
        raise KeyError(name)

    def layer(self, name: str) -> LayerSpec:
        for l in self.layers:
            if l.name == name:
                return l
        raise KeyError(name)

    def names(self) -> List[str]:
        return [t.name for t in self.tensors]

    def min_divisor(self) -> int:
        return 2 ** (1 + len(self.down))

    def check_input(self, h: int, w: int):
        d = self.min_divisor()
        if h % d or w % d:
            raise ValueError(f"input side must be divisible by {d}, got {h}x{w}")

    # -- initialisation (Keras defaults: glorot_uniform kernels, zero bias, BN 1/0/0/1) --
    def init_weights(self, seed: int = 0) -> List[np.ndarray]:
        rng = np.random.default_rng(seed)
        out = []
        for t in self.tensors:
            if t.kind in ("conv_kernel", "pw_kernel"):
                kh, kw, cin, cout = t.shape
                fan_in, fan_out = kh * kw * cin, kh * kw * cout
            elif t.kind == "convT_kernel":
                kh, kw, cout, cin = t.shape
                # Keras computes fans from the stored shape: in = shape[-2], out = shape[-1]
                fan_in, fan_out = kh * kw * cout, kh * kw * cin
            elif t.kind == "dw_kernel":
                kh, kw, cin, mult = t.shape
                fan_in, fan_out = kh * kw * cin, kh * kw * mult
            else:
                fan_in = fan_out = 0
            if fan_in:
                lim = np.sqrt(6.0 / (fan_in + fan_out))
                a = rng.uniform(-lim, lim, size=t.shape).astype(np.float32)
            elif t.kind in ("gamma", "var"):
                a = np.ones(t.shape, np.float32)
            else:
                a = np.zeros(t.shape, np.float32)
            out.append(a)
        return out

    def flops_forward(self, h: int, w: int) -> float:
        """Forward FLOPs per image (2*MAC), as-written graph."""
        macs = 0.0
        hh, ww = h // 2, w // 2
        macs += hh * ww * 27 * self.entry
        prev = self.entry
        for f in self.down:
            macs += hh * ww * (9 * prev + prev * f)
            macs += hh * ww * (9 * f + f * f)
            hh //= 2; ww //= 2
            macs += hh * ww * prev * f
            prev = f
        for f in self.up:
            macs += hh * ww * 9 * prev * f
            macs += hh * ww * 9 * f * f
            hh *= 2; ww *= 2
            macs += hh * ww * prev * f
            prev = f
        macs += hh * ww * prev
        return 2.0 * macs


REFERENCE_SPEC = UNetSpec()
assert REFERENCE_SPEC.num_tensors == 112 and REFERENCE_SPEC.num_params == 2_058_145


def wide_spec(mult: int = 2) -> UNetSpec:
    """North-star extension (BASELINE.json config 4): same graph, every width x ``mult``."""
    return UNetSpec(entry=ENTRY_FILTERS * mult, down=tuple(f * mult for f in DOWN_FILTERS),
                    up=tuple(f * mult for f in UP_FILTERS))
