"""On-device cross-checks of the engine against the fp32 PyTorch oracle (used by tests and
``scripts/check_engine.py``).  Everything here needs a GPU."""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np
import torch

from ..data.generator import synth_crack_batch
from ..models import unet_ref as R
from ..models.spec import REFERENCE_SPEC, UNetSpec


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


def cosine(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def nhwc_to_nchw(t: torch.Tensor) -> torch.Tensor:
    return t.float().permute(0, 3, 1, 2).contiguous()


def oracle_run(spec: UNetSpec, weights: List[np.ndarray], x_u8: np.ndarray, y_u8: np.ndarray, device, loss="bce",
               dtype=torch.float32, bce_clip: bool = False):
    """fp32 oracle forward+backward on `device`.  Returns (loss, acc, taps, grads list, new bn stats)."""
    w = R.to_torch_weights(weights, dtype, device)
    trainable = [i for i, t in enumerate(spec.tensors) if t.trainable]
    for i in trainable:
        w[i].requires_grad_(True)
    x = torch.from_numpy(x_u8).to(device).permute(0, 3, 1, 2).to(dtype) / 255.0
    y = torch.from_numpy(y_u8).to(device).permute(0, 3, 1, 2).to(dtype)
    taps: Dict[str, torch.Tensor] = {}
    logits, stats = R.unet_forward(w, x, training=True, spec=spec, taps=taps, return_logits=True)
    prob = torch.sigmoid(logits)
    l = R.bce_loss(prob, y) if bce_clip else R.bce_logits_loss(logits, y)
    if loss == "bce_dice":
        l = l + R.dice_loss(prob, y)
    l.backward()
    grads = [w[i].grad if w[i].grad is not None else None for i in range(len(w))]
    acc = R.binary_accuracy(prob, y)
    return float(l), float(acc), {k: v.detach() for k, v in taps.items()}, grads, stats, prob.detach()


def engine_vs_oracle(spec: UNetSpec = REFERENCE_SPEC, B=4, H=64, W=64, seed=0, conv_impl="simt", loss="bce",
                     randomize_bn: bool = True) -> dict:
    """Runs ONE forward+backward (no Adam) through the engine and the oracle; returns error tables."""
    from ..models.unet import UNetEngine
    dev = torch.device("cuda")
    weights = spec.init_weights(seed)
    # make BN/bias parameters non-trivial so that every path is exercised
    rng = np.random.default_rng(seed + 1)
    for i, t in enumerate(spec.tensors):
        if not randomize_bn:
            break
        if t.kind == "bias" or t.kind == "beta" or t.kind == "mean":
            weights[i] = rng.normal(0, 0.1, t.shape).astype(np.float32)
        elif t.kind in ("gamma", "var"):
            weights[i] = rng.uniform(0.5, 1.5, t.shape).astype(np.float32)
    x, y = synth_crack_batch(B, H, W, seed)
    eng = UNetEngine(spec, B, H, W, conv_impl=conv_impl, cuda_graph=False, loss=loss)
    eng.load_weights(weights)
    eng.reset_optimizer(); eng.reset_metrics()
    eng.set_batch(x, y)
    eng.zero_region.zero_(); eng._region_dirty = True
    eng._run(eng.grad_prog)  # everything but the optimiser
    torch.cuda.synchronize()
    acc = eng.zero_region[eng.acc_off:eng.acc_off + 6].cpu().numpy()
    eloss, eacc = float(acc[0] / acc[5]), float(acc[1] / acc[5])
    if loss == "bce_dice":
        eloss += 1.0 - float((2 * acc[2] + eng.dice_smooth) / (acc[3] + acc[4] + eng.dice_smooth))
    oloss, oacc, taps, grads, stats, prob = oracle_run(spec, weights, x, y, dev, loss)

    out = {"loss": (eloss, oloss), "acc": (eacc, oacc), "fwd": {}, "grad": {}, "bn_running": {}}
    fmap = {"entry": "x0"}
    for k in range(len(spec.down)):
        fmap[f"down{k}.y1"] = f"e{k}.y1"; fmap[f"down{k}.y2"] = f"e{k}.y2"; fmap[f"down{k}.out"] = f"e{k}.out"
    for j in range(len(spec.up)):
        fmap[f"up{j}.y1"] = f"d{j}.y1"; fmap[f"up{j}.y2"] = f"d{j}.y2"
    for tname, bname in fmap.items():
        out["fwd"][tname] = rel_err(nhwc_to_nchw(eng._bufs[bname]), taps[tname])
    # decoder outputs: oracle is at 2x resolution (upsampled), engine keeps the low-res tensor
    for j in range(len(spec.up)):
        o = taps[f"up{j}.out"][:, :, ::2, ::2]
        out["fwd"][f"up{j}.out(lowres)"] = rel_err(nhwc_to_nchw(eng._bufs[f"d{j}.s"]), o)
    g = eng.grads.detach()
    for i, t in enumerate(spec.tensors):
        if not t.trainable:
            continue
        eg = g[t.offset:t.offset + t.numel].reshape(t.shape)
        og = grads[i]
        if t.kind == "bias" and float(eg.abs().max()) == 0.0 and float(og.abs().max()) < 1e-4:
            out["grad"][t.name] = ("zero-by-BN", float(eg.abs().max()), float(og.abs().max()))
            continue
        out["grad"][t.name] = (rel_err(eg, og), cosine(eg, og), float(og.norm()))
    newa = eng.arena.detach()
    bi = 0
    for l in spec.layers:
        if l.op != "bn":
            continue
        mi, vi = spec.tensors[l.tensors[2]], spec.tensors[l.tensors[3]]
        out["bn_running"][l.name] = (rel_err(newa[mi.offset:mi.offset + mi.numel], stats[bi][0]),
                                     rel_err(newa[vi.offset:vi.offset + vi.numel], stats[bi][1]))
        bi += 1
    return out


def format_report(rep: dict) -> str:
    lines = [f"loss   engine {rep['loss'][0]:.6f}  oracle {rep['loss'][1]:.6f}",
             f"acc    engine {rep['acc'][0]:.6f}  oracle {rep['acc'][1]:.6f}", "-- forward taps (rel L2 err)"]
    for k, v in rep["fwd"].items():
        lines.append(f"  {k:24s} {v:.4e}")
    lines.append("-- gradients (rel L2 err, cosine, |oracle|)")
    for k, v in rep["grad"].items():
        if isinstance(v[0], str):
            lines.append(f"  {k:44s} {v[0]} engine|max|={v[1]:.2e} oracle|max|={v[2]:.2e}")
        else:
            lines.append(f"  {k:44s} {v[0]:.4e}  cos {v[1]:.5f}  |g| {v[2]:.3e}")
    lines.append("-- BN running stats (rel err mean, var)")
    for k, v in rep["bn_running"].items():
        lines.append(f"  {k:28s} {v[0]:.3e} {v[1]:.3e}")
    return "\n".join(lines)


def worst(rep: dict) -> Tuple[float, float]:
    wf = max(rep["fwd"].values())
    wc = min(v[1] for v in rep["grad"].values() if not isinstance(v[0], str))
    return wf, wc


def directional_fd_check(spec: UNetSpec = REFERENCE_SPEC, B=8, H=64, W=64, seed=0, conv_impl="simt",
                         tensors=("conv2d/kernel", "separable_conv2d/depthwise_kernel", "batch_normalization/gamma",
                                  "separable_conv2d_3/pointwise_kernel", "conv2d_transpose/kernel", "conv2d_2/kernel"),
                         rel_step=0.02) -> dict:
    """Finite differences of the ENGINE's own loss along each tensor's engine-gradient direction.

    slope_fd = (L(w + e d) - L(w - e d)) / 2e  with d = g/|g| should equal |g| if the hand-written backward is the
    gradient of the hand-written forward (independent of how chaotic the fp32-vs-bf16 comparison is).
    """
    from ..models.unet import UNetEngine
    weights = spec.init_weights(seed)
    x, y = synth_crack_batch(B, H, W, seed)
    eng = UNetEngine(spec, B, H, W, conv_impl=conv_impl, cuda_graph=False)
    eng.load_weights(weights); eng.reset_optimizer(); eng.set_batch(x, y)
    base = eng.arena.clone()

    def loss_at(flat):
        eng.arena.copy_(flat)
        eng.zero_region.zero_(); eng._region_dirty = True
        eng._run(eng.fwd_prog)
        a = eng.zero_region[eng.acc_off:eng.acc_off + 6].cpu().numpy()
        return float(a[0]) / float(a[5])

    eng.zero_region.zero_(); eng.grads.zero_(); eng._region_dirty = True
    eng._run(eng.grad_prog)
    g = eng.grads.clone()
    _, _, _, ograds, _, _ = oracle_run(spec, weights, x, y, torch.device("cuda"))
    out = {}
    for name in tensors:
        i = spec.index(name)
        t = spec.tensors[i]
        gi = g[t.offset:t.offset + t.numel]
        gn = float(gi.norm())
        d = gi / (gn + 1e-30)
        wn = float(base[t.offset:t.offset + t.numel].norm())
        e = rel_step * max(wn, 1e-3)
        fp, fm = base.clone(), base.clone()
        fp[t.offset:t.offset + t.numel] += e * d
        fm[t.offset:t.offset + t.numel] -= e * d
        slope = (loss_at(fp) - loss_at(fm)) / (2 * e)
        og = ograds[i].reshape(-1)
        out[name] = {"engine_|g|": gn, "fd_slope": slope, "oracle_d.g": float((og * d).sum()), "step": e}
    eng.arena.copy_(base)
    return out


def bf16_oracle_noise(spec: UNetSpec = REFERENCE_SPEC, B=4, H=64, W=64, seed=0) -> dict:
    """How far a plain bf16 PyTorch run of the same graph is from fp32 (context for the engine's errors)."""
    dev = torch.device("cuda")
    weights = spec.init_weights(seed)
    x, y = synth_crack_batch(B, H, W, seed)
    _, _, _, g32, _, _ = oracle_run(spec, weights, x, y, dev)
    _, _, _, g16, _, _ = oracle_run(spec, weights, x, y, dev, dtype=torch.bfloat16)
    return {t.name: cosine(g16[i].float(), g32[i]) for i, t in enumerate(spec.tensors)
            if t.trainable and g32[i] is not None and t.kind != "bias"}
