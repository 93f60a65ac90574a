"""BASELINE config 3 at full size: the four Qwen3-8B shapes at M = 512 ... 8192 (plus a ragged M) through the
tcgen05 prefill kernel -- several m-tiles, several tiles per persistent CTA (gate_up at M = 4096 is 3072 tiles on 148
CTAs), TM = 256 -- the code paths the M = 4096 headline runs on.

Modelled on the reference's own prefill-path test, test_awq_marlin_dequant_matmul_matches_cpu_reference_m32
(marlin_cuda.rs:2474-2612: dequant + GEMM against the scalar oracle, tolerance max(1% of max|y|, 5e-2), :2598-2611).
Two checks per case:
  (a) ALL M x N outputs against  x_f32 @ W_f32  computed in fp32 on the device, W being the reference's D2 grid --
      the reference's own awq_marlin_dequant_int4_bf16 kernel (oracle/_ref) when it is built, else the numpy oracle's
      np_dequant_bf16 (the two are checked bit-identical in test_reference_kernels_gpu.py);
  (b) every row on a fixed random subset of 512 output columns against the float64 CPU oracle (oracle_forward).
Bound for both: max|err| <= 1e-2 * max|y| (the north-star bar) and the reference's tolerance."""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import awq_unpacked, bits_of, t_bf16_from_bits, t_i32

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REF_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_kernels.so")
QWEN3_8B = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 24576), "down": (12288, 4096)}
_cache = {}


def _ref_lib():
    if "ref" not in _cache:
        lib = None
        if os.path.exists(REF_SO):
            lib = ctypes.CDLL(REF_SO)
            vp, i = ctypes.c_void_p, ctypes.c_int
            lib.ref_awq_marlin_dequant_int4_bf16.argtypes = [vp, vp, vp, vp, i, i, i, vp]
        _cache["ref"] = lib
    return _cache["ref"]


def layer(product, name, g, bias):
    """One loaded layer + its dense D2 weight matrix (device, bf16) + the unpacked checkpoint, cached per (shape, g)."""
    key = (name, g, bias)
    if _cache.get("key") != key:
        _cache.pop("val", None)
        torch.cuda.empty_cache()
        K, N = QWEN3_8B[name]
        qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=K ^ N ^ g, bias=bias)
        lin = product.AwqConfig.int4(g).create_linear(K, N, bias, DEV)
        w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
        if bias:
            w["bias"] = torch.from_numpy(b).to(torch.bfloat16)
        lin.load_weights(w)
        q, z, sb = awq_unpacked(qweight, qzeros, scales)
        ref = _ref_lib()
        if ref is not None:   # the reference's own dequant kernel on its own resident layout (awq_marlin.rs:621-656)
            qw_t = t_i32(O.C.transpose_u32(O.C.awq_to_gptq_qweight(qweight, K, N))).to(DEV)
            qz = t_i32(O.C.repack_awq_nibbles(qzeros)).to(DEV)
            sbt = t_bf16_from_bits(sb).to(DEV)
            W = torch.empty((K, N), dtype=torch.bfloat16, device=DEV)
            assert ref.ref_awq_marlin_dequant_int4_bf16(W.data_ptr(), qw_t.data_ptr(), sbt.data_ptr(), qz.data_ptr(), K, N, g,
                                                        torch.cuda.current_stream().cuda_stream) == 0
            torch.cuda.synchronize()
            src = "reference kernel awq_marlin_dequant_int4_bf16"
        else:
            W = t_bf16_from_bits(O.np_dequant_bf16(q, z, sb, g)).to(DEV)
            src = "oracle np_dequant_bf16"
        assert (bits_of(lin.dequantize()) == bits_of(W)).all(), f"product dequant != {src}"
        _cache["key"], _cache["val"] = key, (lin, W, q, z, sb, None if not bias else O.np_bf16_round(b), src)
    return _cache["val"]


def check_case(product, name, g, M, bias):
    K, N = QWEN3_8B[name]
    lin, W, q, z, sb, bias_r, src = layer(product, name, g, bias)
    assert lin.kernel_for(M) == "prefill_umma"
    xb = O.make_activations(M, K, seed=M)
    x = t_bf16_from_bits(xb).to(DEV)
    y = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)     # every element must be written
    lin.forward_into(x, y)
    torch.cuda.synchronize()
    # (a) all outputs, fp32 on the device (no TF32)
    torch.backends.cuda.matmul.allow_tf32 = False
    want = x.float() @ W.float()
    if bias_r is not None:
        want += torch.from_numpy(bias_r).to(DEV)[None, :]
    scale = float(want.abs().max())
    err = float((y.float() - want).abs().max())
    assert np.isfinite(err), f"{name} M={M}: non-finite output"
    assert err <= max(0.01 * scale, 5e-2), f"{name} M={M} g={g}: max err {err} vs reference tolerance (max|y| {scale})"
    assert err <= 1e-2 * scale, f"{name} M={M} g={g}: max err {err} > 1e-2 * max|y| {scale} ({src} + fp32 matmul)"
    del want
    # (b) 512 fixed random columns, all rows, float64 CPU oracle
    cols = np.sort(np.random.default_rng(N + M).choice(N, size=512, replace=False))
    xf = O.np_bf16_bits_to_f32(xb)
    bsub = None if bias_r is None else bias_r[cols]
    want64 = O.np_gemm_ref(xf, np.ascontiguousarray(q[:, cols]), np.ascontiguousarray(z[:, cols]),
                           O.np_bf16_bits_to_f32(np.ascontiguousarray(sb[:, cols])), g, None, bsub, 0)
    got = O.np_bf16_bits_to_f32(bits_of(y[:, torch.from_numpy(cols).to(DEV)])).astype(np.float64)
    err64 = np.abs(got - want64).max()
    assert err64 <= 1e-2 * np.abs(want64).max(), f"{name} M={M}: {err64} vs float64 oracle on 512 columns"
    return err / scale


@pytest.mark.parametrize("M", [512, 1024, 2048, 4096, 4100, 8192])
@pytest.mark.parametrize("name", ["qkv", "o", "gate_up", "down"])   # closest to the function = slowest: one weight set per shape
def test_config3_prefill_full_size(product, name, M):
    check_case(product, name, 128, M, bias=False)


@pytest.mark.parametrize("M", [1024, 4100])
@pytest.mark.parametrize("name", ["qkv", "down"])
def test_config3_prefill_g32_and_bias(product, name, M):
    """Group size 32 (two groups per 64-k stage in the dequant warps) and the bias epilogue at several m-tiles."""
    check_case(product, name, 32, M, bias=True)


def test_many_tiles_per_persistent_cta_is_deterministic(product):
    """gate_up at M = 2048: 192 n-tiles x 8 m-tiles = 1536 tiles on 148 CTAs (10-11 tiles each): twice the same bits."""
    K, N = QWEN3_8B["gate_up"]
    lin, *_ = layer(product, "gate_up", 128, False)
    x = t_bf16_from_bits(O.make_activations(2048, K, seed=3)).to(DEV)
    a = bits_of(lin.forward(x))
    b = bits_of(lin.forward(x))
    assert (a == b).all()
