"""Shared helpers of the parity tests: checkpoint -> product layer, checkpoint -> oracle result."""
from __future__ import annotations

import numpy as np
import torch

from oracle import oracle as O


def t_i32(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int32))


def t_bf16_from_bits(bits: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(bits).view(np.int16)).view(torch.bfloat16)


def bits_of(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def awq_unpacked(qweight, qzeros, scales):
    """-> q [K,N] u8, z [G,N] u8, bf16 scale bits [G,N] (S1: f16 -> bf16 RNE)."""
    q, z = O.np_awq_unpack(qweight), O.np_awq_unpack(qzeros)
    sb = O.np_f32_to_bf16_bits(np.asarray(scales, np.float32))
    return q, z, sb


def gptq_unpacked(qweight, qzeros, scales):
    q = O.np_gptq_unpack_qweight(qweight)
    z = O.np_seq_unpack(qzeros) if qzeros is not None else None
    sb = O.np_f32_to_bf16_bits(np.asarray(scales, np.float32))
    return q, z, sb


def oracle_forward(x_bits, q, z, sb, group_size, g_idx=None, bias_f32=None, zero_mode=0, exact_small=False):
    """fp32 dequant-matmul on the bf16-rounded operands (marlin_cuda.rs:2083-2112).
    exact_small=True runs the scalar C loop (sequential fp32, the reference's own order)."""
    x = O.np_bf16_bits_to_f32(x_bits)
    s = O.np_bf16_bits_to_f32(sb)
    K, N = q.shape
    if zero_mode == 1 and z is None:
        z = np.full((s.shape[0], N), 8, np.uint8)
    if exact_small:
        gi = None if g_idx is None else np.ascontiguousarray(g_idx, np.int32)
        b = None if bias_f32 is None else np.ascontiguousarray(bias_f32, np.float32)
        return O.C.gemm_f32(np.ascontiguousarray(x), q, z, np.ascontiguousarray(s), gi, b,
                            group_size if group_size > 0 else 0, K, N, zero_mode, threads=O.C.max_threads())
    return O.np_gemm_ref(x, q, z, s, group_size, g_idx, bias_f32, zero_mode)


def assert_close_ref(y_bits, expected, what=""):
    """The reference's own tolerance, max(0.01*max|exp|, 5e-2) (marlin_cuda.rs:2168-2183), AND the
    north-star bound: max |err| <= 1e-2 * max|exp| (bf16 output vs fp32 dequant-matmul)."""
    got = O.np_bf16_bits_to_f32(y_bits).astype(np.float64)
    err = np.abs(got - expected).max()
    scale = float(np.abs(expected).max())
    assert err <= O.reference_tolerance(expected), f"{what}: max err {err} > reference tolerance"
    assert err <= 1e-2 * max(scale, 1e-6) + 1e-6, f"{what}: max err {err} vs max|y| {scale}: rel > 1e-2"
    return err, scale
