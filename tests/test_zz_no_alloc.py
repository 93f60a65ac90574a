"""w4a16_forward_decode_static: the decode forward whose split-K scratch is a static device array (no cudaMalloc).

No counterpart in the reference.  Collected last on purpose.
The check is the existing entry point, w4a16_forward_into, on the same handle and inputs, within the
GEMM tolerance the parity tests use (1e-2 * max|y|), on two calls in a row (the array keeps the last call's contents).
"""
import ctypes

import pytest

import conftest

pytestmark = pytest.mark.gpu

TOL = 1e-2   # of max|y|, as tests/test_parity_gpu.py


def _linear(P, torch, K, N, g, seed):
    gen = torch.Generator(device="cpu").manual_seed(seed)
    qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, generator=gen)
    qzeros = torch.randint(-2**31, 2**31 - 1, (K // g, N // 8), dtype=torch.int32, generator=gen)
    scales = (torch.rand((K // g, N), generator=gen) * 0.018 + 0.002).to(torch.float16)
    lin = P.AwqConfig.int4(g).create_linear(K, N, False, "cuda:0")
    lin.load_weights({"qweight": qweight.cuda(), "qzeros": qzeros.cuda(), "scales": scales.cuda()})
    return lin


def _static(P, lin, x, y, M):
    import torch
    return P.lib().w4a16_forward_decode_static(lin._h, ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(y.data_ptr()), M,
                                               ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))


@pytest.mark.parametrize("K,N,g,M", [(4096, 4096, 128, 8), (4096, 4096, 128, 1), (12288, 4096, 128, 16), (2048, 1024, 64, 27)])
def test_static_scratch_decode_matches_forward_into(K, N, g, M):
    import torch
    P = conftest._load_product()
    lin = _linear(P, torch, K, N, g, seed=K + N + M)
    gen = torch.Generator(device="cpu").manual_seed(M)
    for call in range(2):   # two calls in a row: the second starts from the first one's tickets and partials
        x = torch.randn((M, K), generator=gen).to(torch.bfloat16).cuda()
        want = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device="cuda:0")
        got = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device="cuda:0")
        lin.forward_into(x, want)
        rc = _static(P, lin, x, got, M)
        assert rc == 0, P.lib().w4a16_last_error().decode()
        torch.cuda.synchronize()
        scale = want.float().abs().max().item()
        err = (got.float() - want.float()).abs().max().item()
        assert torch.isfinite(got.float()).all(), f"call {call}: non-finite output"
        assert err <= TOL * scale, f"call {call}: max|err| {err} > {TOL} * {scale}"


def test_static_scratch_decode_rejects_what_it_cannot_hold():
    import torch
    P = conftest._load_product()
    lin = _linear(P, torch, 1024, 1024, 128, seed=1)
    x = torch.zeros((64, 1024), dtype=torch.bfloat16, device="cuda:0")
    y = torch.zeros((64, 1024), dtype=torch.bfloat16, device="cuda:0")
    assert _static(P, lin, x, y, 33) != 0
    assert "M must be in 1..32" in P.lib().w4a16_last_error().decode()
    assert _static(P, lin, x, y, 0) != 0
    assert P.lib().w4a16_forward_decode_static(lin._h, None, ctypes.c_void_p(y.data_ptr()), 1, None) != 0
    assert "null x or y" in P.lib().w4a16_last_error().decode()
