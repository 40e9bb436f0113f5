"""What `--use_fast_math` changes.  The library is built with it (approximate division, sqrtf, logf, expf); the same
optimiser / elementwise sources are also built without it (crackfl/_lib/libcrackfl_ieee.so) and the kernels whose
arithmetic it touches - Keras-form Adam, BN finalize, the BCE head - are run from both builds on identical inputs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _st():
    return torch.cuda.current_stream().cuda_stream


def test_fast_math_build_stays_within_a_few_ulp_of_the_ieee_build():
    from crackfl.ops import lib as L
    fast, ieee = L.get_lib(), L.get_ieee_lib()
    torch.manual_seed(0)
    # ---- Adam: theta -= lr_t * m / (sqrt(v) + eps), 5 steps (client_fit_model.py:157)
    n = 1 << 16
    res = {}
    for name, lib in (("fast", fast), ("ieee", ieee)):
        g0 = torch.Generator(device="cuda").manual_seed(1)
        p = torch.randn(n, device="cuda", generator=g0)
        m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
        step = torch.zeros(1, dtype=torch.int32, device="cuda")
        for it in range(5):
            g = torch.randn(n, device="cuda", generator=g0) * (10.0 ** (it - 4))  # gradients from 1e-4 to 1
            L.check(lib.crackfl_adam_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), step.data_ptr(), n, 1e-3,
                                          0.9, 0.999, 1e-7, _st()))
        torch.cuda.synchronize()
        res[name] = p.clone()
    upd = (res["ieee"] - res["fast"]).abs().max().item()
    # measured 7e-7: ~1.5 fp32 ulp of a parameter of magnitude 2-4, and 0.014 % of the 5e-3 the five steps moved it
    assert upd < 2e-6, upd
    # ---- BN finalize: scale = gamma * rsqrt(var + eps), running statistics
    C_, P = 256, 65536.0
    for_bn = {}
    for name, lib in (("fast", fast), ("ieee", ieee)):
        g0 = torch.Generator(device="cuda").manual_seed(2)
        mean = torch.randn(C_, device="cuda", generator=g0)
        var = torch.rand(C_, device="cuda", generator=g0) * 4 + 1e-4
        stats = torch.cat([mean * P, (var + mean * mean) * P])
        gamma = torch.rand(C_, device="cuda", generator=g0) + 0.5; beta = torch.randn(C_, device="cuda", generator=g0)
        rm = torch.zeros(C_, device="cuda"); rv = torch.ones(C_, device="cuda")
        bn = torch.zeros(4 * C_, device="cuda")
        L.check(lib.crackfl_bn_finalize(stats.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                                        bn.data_ptr(), C_, P, 1e-3, 0.99, _st()))
        torch.cuda.synchronize()
        for_bn[name] = (bn.clone(), rm.clone(), rv.clone())
    for a, b in zip(for_bn["fast"], for_bn["ieee"]):
        assert torch.allclose(a, b, rtol=2e-6, atol=1e-6), float((a - b).abs().max())
    # ---- BCE head: loss sums and the gradient written for the decoder (logits from -30 to 30)
    B, Hl, Wl, Cc = 2, 16, 16, 32
    out = {}
    for name, lib in (("fast", fast), ("ieee", ieee)):
        g0 = torch.Generator(device="cuda").manual_seed(3)
        s = (torch.randn(B, Hl, Wl, Cc, device="cuda", generator=g0) * 3).bfloat16()
        w = torch.randn(Cc, device="cuda", generator=g0); bias = torch.zeros(1, device="cuda")
        mask = (torch.rand(B, 2 * Hl, 2 * Wl, device="cuda", generator=g0) > 0.9).to(torch.uint8)
        acc = torch.zeros(8, device="cuda"); ds = torch.zeros_like(s)
        dw = torch.zeros(Cc, device="cuda"); db = torch.zeros(1, device="cuda")
        L.check(lib.crackfl_head_fwd_bwd(s.data_ptr(), w.data_ptr(), bias.data_ptr(), mask.data_ptr(), acc.data_ptr(),
                                         ds.data_ptr(), dw.data_ptr(), db.data_ptr(), B, Hl, Wl, Cc, 1.0, 0, _st()))
        torch.cuda.synchronize()
        out[name] = (acc.clone(), ds.float().clone(), dw.clone())
    loss_f, loss_i = out["fast"][0][0] / out["fast"][0][5], out["ieee"][0][0] / out["ieee"][0][5]
    assert abs(float(loss_f - loss_i)) < 1e-5 * max(1.0, float(loss_i)), (float(loss_f), float(loss_i))
    assert float((out["fast"][1] - out["ieee"][1]).abs().max()) <= 2 ** -8 * float(out["ieee"][1].abs().max())  # bf16 ulp
    assert torch.allclose(out["fast"][2], out["ieee"][2], rtol=1e-4, atol=1e-6)
