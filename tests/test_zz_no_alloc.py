"""Allocation-free BatchNorm forward (`crackfl_bn_train_fwd_noalloc`) through the C ABI.

The entry point exists for users of libcrackfl.so that are not in a position to allocate device memory
themselves - a plug-in whose host process has given the whole device heap to another framework, a step function
running under a fixed memory budget, error-recovery code after an out-of-memory failure.  Its scratch is a
fixed array inside the library, shared by all calls on a GPU, so this test checks the two properties such a
caller relies on: the result equals the ordinary path (which takes caller-provided scratch), and a second call
right after the first - with different data and a different channel count, so stale scratch would show - is
just as correct.  The file name sorts last on purpose: a failure here never hides another test's result.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference_path(L, lib, y, gamma, beta, rm, rv, B, H, W, Cc, relu, eps, mom):
    """Ordinary entry points with scratch owned by the caller (here: torch tensors)."""
    st = torch.cuda.current_stream().cuda_stream
    stats = torch.zeros(2 * Cc, device="cuda")
    bn = torch.zeros(4 * Cc, device="cuda")
    out = torch.empty_like(y)
    L.check(lib.crackfl_bn_stats(y.data_ptr(), stats.data_ptr(), B * H * W, Cc, st))
    L.check(lib.crackfl_bn_finalize(stats.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                                    bn.data_ptr(), Cc, float(B * H * W), eps, mom, st))
    L.check(lib.crackfl_bn_act(y.data_ptr(), bn.data_ptr(), out.data_ptr(), 0, B, H, W, Cc, relu, st))
    return out


def _case(L, lib, seed, B, H, W, Cc, relu):
    torch.manual_seed(seed)
    y = (torch.randn(B, H, W, Cc, device="cuda") * 2 + 0.5).bfloat16()
    gamma = torch.rand(Cc, device="cuda") + 0.5
    beta = torch.randn(Cc, device="cuda") * 0.1
    rm_a, rv_a = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
    rm_b, rv_b = rm_a.clone(), rv_a.clone()
    eps, mom = 1e-3, 0.99
    want = _reference_path(L, lib, y, gamma, beta, rm_a, rv_a, B, H, W, Cc, relu, eps, mom)
    got = torch.empty_like(y)
    rc = lib.crackfl_bn_train_fwd_noalloc(y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm_b.data_ptr(),
                                          rv_b.data_ptr(), got.data_ptr(), B, H, W, Cc, relu, eps, mom,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    err = float((got.float() - want.float()).norm() / want.float().norm())
    assert err < 1e-2, err
    assert torch.allclose(rm_a, rm_b, rtol=1e-4, atol=1e-6) and torch.allclose(rv_a, rv_b, rtol=1e-4, atol=1e-6)
    # and both agree with plain PyTorch
    yf = y.float().reshape(-1, Cc)
    ref = (yf - yf.mean(0)) * torch.rsqrt(yf.var(0, unbiased=False) + eps) * gamma + beta
    if relu:
        ref = ref.relu()
    assert float((got.float().reshape(-1, Cc) - ref).norm() / ref.norm()) < 1e-2


def test_noalloc_bn_matches_ordinary_path_twice_in_a_row():
    from crackfl.ops import lib as L
    lib = L.get_lib()
    assert lib.crackfl_noalloc_max_channels() == 2048
    _case(L, lib, seed=0, B=4, H=16, W=16, Cc=256, relu=1)
    _case(L, lib, seed=1, B=2, H=24, W=8, Cc=64, relu=0)   # fewer channels: leftovers of call 1 stay in the array


def test_noalloc_bn_rejects_oversized_channel_count():
    from crackfl.ops import lib as L
    lib = L.get_lib()
    y = torch.zeros(8, device="cuda", dtype=torch.bfloat16)
    f = torch.zeros(8, device="cuda")
    rc = lib.crackfl_bn_train_fwd_noalloc(y.data_ptr(), f.data_ptr(), f.data_ptr(), f.data_ptr(), f.data_ptr(),
                                          y.data_ptr(), 1, 1, 1, 4096, 0, 1e-3, 0.99,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc != 0
