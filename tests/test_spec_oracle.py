"""Model spec + Keras-semantics oracle (SURVEY §2.5, §4 'Unit, CPU')."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from crackfl.models.spec import REFERENCE_SPEC, UNetSpec, wide_spec
from crackfl.models import unet_ref as R


def test_reference_spec_counts():
    s = REFERENCE_SPEC
    assert s.num_tensors == 112
    assert s.num_params == 2_058_145
    assert s.num_trainable == 2_054_369
    assert s.tensors[0].name == "conv2d/kernel:0" and s.tensors[0].shape == (3, 3, 3, 32)
    assert s.tensors[-1].name == "conv2d_8/bias:0"
    # block totals from SURVEY §2.5
    by_layer = {}
    for t in s.tensors:
        by_layer[t.layer] = by_layer.get(t.layer, 0) + t.numel
    assert by_layer["conv2d_transpose"] == 590_080 and by_layer["separable_conv2d"] == 2_400
    assert abs(s.flops_forward(128, 128) / 1e9 - 1.249) < 0.01
    assert abs(s.flops_forward(256, 256) / 1e9 - 4.998) < 0.01
    assert len(set(s.names())) == 112


def test_wide_spec():
    w = wide_spec(2)
    assert w.num_tensors == 112 and w.num_params > 4 * REFERENCE_SPEC.num_params * 0.9


def test_forward_shapes_and_padding():
    s = REFERENCE_SPEC
    w = R.to_torch_weights(s.init_weights(0))
    x = torch.rand(2, 3, 64, 64)
    p, stats = R.unet_forward(w, x, training=True)
    assert p.shape == (2, 1, 64, 64) and len(stats) == 15
    assert float(p.min()) >= 0 and float(p.max()) <= 1
    with pytest.raises(ValueError):
        R.unet_forward(w, torch.rand(1, 3, 40, 40))


def test_tf_same_padding_is_bottom_right():
    x = torch.arange(16.0).reshape(1, 1, 4, 4)
    p = R._same_pad_s2(x)
    assert p.shape[-2:] == (5, 5)
    assert torch.equal(p[0, 0, :4, :4], x[0, 0]) and float(p[0, 0, 4].abs().sum()) == 0


def test_convT_equals_flipped_conv():
    """Conv2DTranspose(stride 1, same) == 3x3 conv with flipped taps and swapped channels."""
    k = torch.randn(3, 3, 5, 4)  # (kh,kw,out,in)
    x = torch.randn(2, 4, 8, 8)
    a = F.conv_transpose2d(x, k.permute(3, 2, 0, 1), None, stride=1, padding=1)
    wflip = k.flip(0, 1).permute(2, 3, 0, 1)  # (out,in,kh,kw)
    b = F.conv2d(x, wflip, None, padding=1)
    assert torch.allclose(a, b, atol=1e-5)


def test_residual_commutes_with_upsample():
    x = torch.randn(1, 6, 5, 5); w = torch.randn(3, 6, 1, 1)
    a = F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), w)
    b = F.interpolate(F.conv2d(x, w), scale_factor=2, mode="nearest")
    assert torch.allclose(a, b, atol=1e-6)


def test_keras_adam_matches_formula():
    p = torch.tensor([1.0, -2.0]); g = torch.tensor([0.5, 0.25])
    opt = R.KerasAdam([p])
    opt.step([g])
    m = 0.1 * g; v = 0.001 * g * g
    lr_t = 1e-3 * np.sqrt(1 - 0.999) / (1 - 0.9)
    assert torch.allclose(p, torch.tensor([1.0, -2.0]) - lr_t * m / (v.sqrt() + 1e-7), atol=1e-7)


def test_bce_clip():
    p = torch.tensor([0.0, 1.0]); y = torch.tensor([1.0, 0.0])
    # fp32: 1-1e-7 rounds to 1-2^-23, so the clipped loss is ~16.0 (same in Keras fp32)
    assert abs(float(R.bce_loss(p, y)) - (-np.log(1e-7))) < 0.2


def test_bce_from_logits_is_the_keras2_fit_form():
    """sigmoid_cross_entropy_with_logits: same value as the probability form where that is finite, and a gradient
    sigmoid(z)-y that does not vanish where the clipped form has none."""
    z = torch.tensor([-30.0, -3.0, 0.5, 4.0, 30.0], requires_grad=True)
    y = torch.tensor([1.0, 0.0, 1.0, 0.0, 0.0])
    l = R.bce_logits_loss(z, y)
    l.backward()
    assert torch.allclose(z.grad, (torch.sigmoid(z.detach()) - y) / 5, atol=1e-7)
    mid = slice(1, 4)
    ref = F.binary_cross_entropy(torch.sigmoid(z.detach()[mid]), y[mid])
    assert abs(float(R.bce_logits_loss(z.detach()[mid], y[mid])) - float(ref)) < 1e-6
    zc = z.detach().clone().requires_grad_(True)
    R.bce_loss(torch.sigmoid(zc), y).backward()
    assert float(zc.grad[0]) == 0.0 and float(z.grad[0]) < -0.19  # the clipped form is dead at z = -30, y = 1


def test_ref_trainer_loss_decreases_and_bn_updates():
    from crackfl.data.generator import synth_crack_batch
    s = REFERENCE_SPEC
    tr = R.RefTrainer(s.init_weights(1), s)
    xs, ys = synth_crack_batch(4, 32, 32, 0)
    x = torch.from_numpy(xs).permute(0, 3, 1, 2).float() / 255
    y = torch.from_numpy(ys).permute(0, 3, 1, 2).float()
    losses = [tr.train_step(x, y)[0] for _ in range(8)]
    assert losses[-1] < losses[0]
    w = tr.get_weights()
    mm = w[s.index("batch_normalization/moving_mean")]
    assert np.abs(mm).sum() > 0  # moving stats moved off zero
