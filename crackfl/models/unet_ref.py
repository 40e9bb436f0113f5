"""Keras-semantics U-Net restated in plain PyTorch (fp32/fp64, any device).

This is the numerical ORACLE for every sm_100a kernel and the CPU back-end used by the
plumbing configuration (BASELINE.json config 1, "2-client FedAvg ... on CPU").  It is a
restatement of the graph at ``client_fit_model.py:92-150`` with the Keras/TF details
that differ from PyTorch defaults made explicit (hard part H1):

* TF "same" padding for stride 2 / k=3 on even sizes pads bottom/right only;
* ``Conv2DTranspose(f, 3, padding="same")`` stride 1, kernel stored ``(kh,kw,out,in)``;
* BatchNormalization eps=1e-3, momentum=0.99; running variance gets the unbiased batch
  variance (TF fused batch norm behaviour), normalisation uses the biased one;
* binary cross-entropy: Keras 2 ``fit`` on a model whose output layer carries ``activation="sigmoid"``
  (``client_fit_model.py:145,157``) does NOT use the clipped-probability formula: ``backend.binary_crossentropy``
  recovers the sigmoid's input (``_keras_logits`` / the op type ``Sigmoid``) and calls
  ``tf.nn.sigmoid_cross_entropy_with_logits``, whose gradient ``sigmoid(z) - y`` never vanishes.  That is the default
  here (``bce_logits_loss``).  The clipped form (``bce_loss``: probabilities in [1e-7, 1-1e-7], zero gradient outside,
  the backend's fall-back and Keras 3's formula) is kept behind ``bce_clip=True``; with it a model whose logits all
  pass -16.1 stops training for good (measured: ``scripts/fed_convergence.py``);
* Keras Adam: ``theta -= lr_t * m / (sqrt(v) + eps)``, eps=1e-7 outside the bias correction.

The weights argument is always the list of 112 tensors in ``UNetSpec`` order/layout.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from .spec import REFERENCE_SPEC, UNetSpec

BCE_EPS = 1e-7  # keras.backend.epsilon()


def to_torch_weights(weights: Sequence, dtype=torch.float32, device="cpu", requires_grad=False):
    out = []
    for w in weights:
        t = torch.as_tensor(np.asarray(w) if not torch.is_tensor(w) else w).to(device=device, dtype=dtype).clone()
        out.append(t)
    if requires_grad:
        for t in out:
            t.requires_grad_(True)
    return out


def _same_pad_s2(x, value=0.0):
    # TF SAME, k=3, s=2: total pad = max((ceil(n/2)-1)*2+3-n, 0); extra goes bottom/right.
    h, w = x.shape[-2:]
    ph = max((-(-h // 2) - 1) * 2 + 3 - h, 0)
    pw = max((-(-w // 2) - 1) * 2 + 3 - w, 0)
    return F.pad(x, (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2), value=value)


def _bn(x, gamma, beta, mean, var, training: bool, eps: float, momentum: float, new_stats: list):
    if training:
        n = x.numel() // x.shape[1]
        bm = x.mean(dim=(0, 2, 3))
        bv = x.var(dim=(0, 2, 3), unbiased=False)
        with torch.no_grad():
            unbiased = bv * (n / max(n - 1, 1))
            new_stats.append((mean * momentum + bm.detach() * (1 - momentum),
                              var * momentum + unbiased.detach() * (1 - momentum)))
        m, v = bm, bv
    else:
        new_stats.append((mean, var))
        m, v = mean, var
    inv = torch.rsqrt(v + eps)
    return (x - m[None, :, None, None]) * (inv * gamma)[None, :, None, None] + beta[None, :, None, None]


def unet_forward(weights: Sequence[torch.Tensor], x: torch.Tensor, training: bool = False,
                 spec: UNetSpec = REFERENCE_SPEC, return_logits: bool = False,
                 taps: Optional[Dict[str, torch.Tensor]] = None):
    """Forward pass.  ``x`` is NCHW in [0,1].  Returns ``(prob_or_logits NCHW, new_bn_stats)``.

    ``new_bn_stats`` is a list of ``(moving_mean, moving_variance)`` per BN layer in
    graph order (equal to the inputs when ``training`` is False).  ``taps`` (optional
    dict) receives named intermediate activations for kernel-level tests.
    """
    spec.check_input(x.shape[-2], x.shape[-1])
    it = iter(weights)
    nxt = lambda: next(it)
    stats: list = []
    eps, mom = spec.bn_eps, spec.bn_momentum

    def tap(name, t):
        if taps is not None:
            taps[name] = t
        return t

    def conv(x, stride, k):
        kern, bias = nxt(), nxt()  # (kh,kw,in,out)
        w = kern.permute(3, 2, 0, 1)
        if k == 3 and stride == 2:
            x = _same_pad_s2(x)
            return F.conv2d(x, w, bias, stride=2)
        if k == 1:
            return F.conv2d(x, w, bias, stride=stride)
        return F.conv2d(x, w, bias, stride=1, padding=k // 2)

    def bn(x):
        g, b, m, v = nxt(), nxt(), nxt(), nxt()
        return _bn(x, g, b, m, v, training, eps, mom, stats)

    def sep(x):
        dw, pw, bias = nxt(), nxt(), nxt()
        c = x.shape[1]
        x = F.conv2d(x, dw.permute(2, 3, 0, 1), None, padding=1, groups=c)
        return F.conv2d(x, pw.permute(3, 2, 0, 1), bias)

    def convT(x):
        kern, bias = nxt(), nxt()  # (kh,kw,out,in)
        return F.conv_transpose2d(x, kern.permute(3, 2, 0, 1), bias, stride=1, padding=1)

    x = tap("entry", F.relu(bn(conv(x, 2, 3))))
    prev = x
    for bi, _f in enumerate(spec.down):
        x = F.relu(x)
        x = tap(f"down{bi}.bn1", bn(tap(f"down{bi}.y1", sep(x))))
        x = F.relu(x)
        x = tap(f"down{bi}.bn2", bn(tap(f"down{bi}.y2", sep(x))))
        x = F.max_pool2d(_same_pad_s2(x, float("-inf")), 3, 2)
        res = conv(prev, 2, 1)
        x = tap(f"down{bi}.out", x + res)
        prev = x
    for bi, _f in enumerate(spec.up):
        x = F.relu(x)
        x = tap(f"up{bi}.bn1", bn(tap(f"up{bi}.y1", convT(x))))
        x = F.relu(x)
        x = tap(f"up{bi}.bn2", bn(tap(f"up{bi}.y2", convT(x))))
        x = F.interpolate(x, scale_factor=2, mode="nearest")
        res = conv(F.interpolate(prev, scale_factor=2, mode="nearest"), 1, 1)
        x = tap(f"up{bi}.out", x + res)
        prev = x
    logits = tap("logits", conv(x, 1, 1))
    return (logits if return_logits else torch.sigmoid(logits)), stats


def bce_loss(prob: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """Keras ``binary_crossentropy`` on probabilities, mean over every pixel."""
    p = prob.clamp(BCE_EPS, 1.0 - BCE_EPS)
    return -(y * torch.log(p) + (1.0 - y) * torch.log(1.0 - p)).mean()


def bce_logits_loss(logits: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """``tf.nn.sigmoid_cross_entropy_with_logits`` averaged over every pixel: max(z,0) - z*y + log(1+exp(-|z|))."""
    return (logits.clamp_min(0) - logits * y + torch.log1p(torch.exp(-logits.abs()))).mean()


def dice_loss(prob: torch.Tensor, y: torch.Tensor, smooth: float = 1.0) -> torch.Tensor:
    """Soft Dice loss over the whole batch (north-star extension; not in the reference)."""
    inter = (prob * y).sum()
    return 1.0 - (2.0 * inter + smooth) / (prob.sum() + y.sum() + smooth)


def binary_accuracy(prob: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """Keras ``accuracy`` for a sigmoid output resolves to binary accuracy at 0.5."""
    return ((prob > 0.5).to(y.dtype) == y).to(torch.float32).mean()


class KerasAdam:
    """Keras/TF2 Adam (``client_fit_model.py:157`` ``optimizer="Adam"``): lr 1e-3, b1 .9, b2 .999, eps 1e-7."""

    def __init__(self, params: Sequence[torch.Tensor], lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7):
        self.params = list(params)
        self.lr, self.b1, self.b2, self.eps = lr, beta1, beta2, eps
        self.t = 0
        self.m = [torch.zeros_like(p) for p in self.params]
        self.v = [torch.zeros_like(p) for p in self.params]

    @torch.no_grad()
    def step(self, grads: Sequence[Optional[torch.Tensor]]):
        self.t += 1
        lr_t = self.lr * np.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
        for p, g, m, v in zip(self.params, grads, self.m, self.v):
            if g is None:
                continue
            m.mul_(self.b1).add_(g, alpha=1 - self.b1)
            v.mul_(self.b2).addcmul_(g, g, value=1 - self.b2)
            p.sub_(lr_t * m / (v.sqrt() + self.eps))


class RefTrainer:
    """Reference-semantics local trainer on the oracle (CPU plumbing back-end).

    One instance == one ``compile()``: fresh Adam state, as the reference rebuilds the
    model and optimizer every round (``client_fit_model.py:155-157``).
    """

    def __init__(self, weights: Sequence, spec: UNetSpec = REFERENCE_SPEC, lr: float = 1e-3,
                 loss: str = "bce", dtype=torch.float32, device="cpu", bce_clip: bool = False):
        self.spec = spec
        self.bce_clip = bce_clip
        self.w = to_torch_weights(weights, dtype, device)
        self.trainable = [i for i, t in enumerate(spec.tensors) if t.trainable]
        for i in self.trainable:
            self.w[i].requires_grad_(True)
        self.opt = KerasAdam([self.w[i] for i in self.trainable], lr=lr)
        self.loss_kind = loss
        self.bn_slots = [(l.tensors[2], l.tensors[3]) for l in spec.layers if l.op == "bn"]

    def _loss(self, logits, y):
        prob = torch.sigmoid(logits)
        l = bce_loss(prob, y) if self.bce_clip else bce_logits_loss(logits, y)
        if self.loss_kind == "bce_dice":
            l = l + dice_loss(prob, y)
        return l, prob

    def train_step(self, x: torch.Tensor, y: torch.Tensor) -> Tuple[float, float]:
        """x NCHW float in [0,1], y N1HW in {0,1}.  Returns (loss, accuracy)."""
        for i in self.trainable:
            self.w[i].grad = None
        logits, stats = unet_forward(self.w, x, training=True, spec=self.spec, return_logits=True)
        loss, prob = self._loss(logits, y)
        loss.backward()
        self.opt.step([self.w[i].grad for i in self.trainable])
        with torch.no_grad():
            for (mi, vi), (m, v) in zip(self.bn_slots, stats):
                self.w[mi].copy_(m)
                self.w[vi].copy_(v)
            acc = binary_accuracy(prob, y)
        return float(loss.detach()), float(acc)

    @torch.no_grad()
    def eval_step(self, x, y) -> Tuple[float, float]:
        logits, _ = unet_forward(self.w, x, training=False, spec=self.spec, return_logits=True)
        loss, prob = self._loss(logits, y)
        return float(loss), float(binary_accuracy(prob, y))

    @torch.no_grad()
    def predict(self, x) -> torch.Tensor:
        return unet_forward(self.w, x, training=False, spec=self.spec)[0]

    def get_weights(self) -> List[np.ndarray]:
        return [w.detach().cpu().to(torch.float32).numpy().copy() for w in self.w]
