"""sibling projections fused at load (w4a16_load_*_part) and the SwiGLU epilogue.

The reference runs q/k/v and gate/up as separate quantized linears followed by silu_and_mul_separate_bf16 /
fused_swiglu_bf16 (qwen3_quantized.rs:84-96, 179-181, 532-551; activations.cu:148-164; swiglu.cu:40-45).  Checks:
  * a handle filled part by part is the SAME resident layout as a handle loaded from the column-wise concatenated
    checkpoint (unpack / scales / dequant hooks and forward outputs bit for bit), AWQ and GPTQ, bias, act-order;
  * each part of the fused output against the float64 oracle of that sibling alone (the reference tolerance);
  * the SwiGLU handle against the REFERENCE'S OWN kernels (oracle/_ref, both of them) applied to the outputs of two
    separate product linears -- bit for bit on inputs whose fp32 sums are exact in any order (small-integer x,
    power-of-two scales), so the only arithmetic under test is the epilogue's -- and against the numpy oracle
    (np_swiglu_bf16) within one bf16 ulp; on N(0,1) inputs within the north-star tolerance;
  * the error behaviour of the part loads."""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import assert_close_ref, awq_unpacked, bits_of, gptq_unpacked, oracle_forward, t_bf16_from_bits, t_i32

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REF_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_kernels.so")
M_LIST = [1, 2, 3, 4, 7, 16, 33, 64, 130, 300]


def _ref():
    if not os.path.exists(REF_SO):
        return None
    lib = ctypes.CDLL(REF_SO)
    vp, i = ctypes.c_void_p, ctypes.c_int
    for name in ("ref_fused_swiglu_bf16", "ref_silu_and_mul_separate_bf16"):
        if not hasattr(lib, name):
            return None
        getattr(lib, name).argtypes = [vp, vp, vp, i, i, vp]
    return lib


def awq_weights(K, N, g, seed, bias=False):
    qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=seed, bias=bias)
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
    if bias:
        w["bias"] = torch.from_numpy(b).to(torch.bfloat16)
    return w, (qweight, qzeros, scales, b)


def gptq_weights(K, N, g, seed, g_idx=None):
    qweight, qzeros, scales, _ = O.make_gptq_checkpoint(K, N, g, seed=seed)
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
    if g_idx is not None:
        w["g_idx"] = torch.from_numpy(g_idx)
    return w, (qweight, qzeros, scales)


def cat_awq(ws):
    """Column-wise concatenation of AWQ checkpoints: every tensor is [*, N/8] or [*, N] -> cat on the last dim."""
    out = {k: torch.cat([w[k] for w in ws], dim=-1) for k in ws[0]}
    return out


def same_handle(a, b):
    qa, za = a.unpack(); qb, zb = b.unpack()
    assert torch.equal(qa, qb) and torch.equal(za, zb)
    assert (bits_of(a.scales()) == bits_of(b.scales())).all()
    assert (bits_of(a.dequantize()) == bits_of(b.dequantize())).all()


# ---- fused parts == concatenated checkpoint ----------------------------------------------------------------------------
@pytest.mark.parametrize("K,parts,g,bias", [(4096, (4096, 1024, 1024), 128, False),   # Qwen3-8B qkv -> imma2
                                            (2048, (2048, 1024, 1024), 128, True),    # Qwen3-0.6B-like qkv, bias -> imma v1
                                            (1024, (256, 64, 64), 32, True),
                                            (512, (48, 16), 64, False)])              # N = 64: the stream kernel
def test_awq_parts_equal_concatenated_checkpoint(product, K, parts, g, bias):
    ws = [awq_weights(K, n, g, seed=100 + i, bias=bias)[0] for i, n in enumerate(parts)]
    N = sum(parts)
    cfg = product.AwqConfig.int4(g)
    whole = cfg.create_linear(K, N, bias, DEV)
    whole.load_weights(cat_awq(ws))
    fused = cfg.create_linear(K, N, bias, DEV)
    for i in ((1, 0, 2) if len(parts) == 3 else (1, 0)):     # any order
        fused.load_part(sum(parts[:i]), ws[i])
    same_handle(fused, whole)
    for M in M_LIST:
        x = t_bf16_from_bits(O.make_activations(M, K, seed=M)).to(DEV)
        assert fused.kernel_for(M) == whole.kernel_for(M)
        assert (bits_of(fused.forward(x)) == bits_of(whole.forward(x))).all(), f"M={M}"


def test_fused_parts_against_the_oracle_of_each_sibling(product):
    K, parts, g = 2048, (512, 256, 256), 128
    cks = [awq_weights(K, n, g, seed=7 + i, bias=True) for i, n in enumerate(parts)]
    fused = product.AwqConfig.int4(g).create_linear(K, sum(parts), True, DEV)
    n0 = 0
    for (w, _), n in zip(cks, parts):
        fused.load_part(n0, w)
        n0 += n
    for M in (1, 4, 16, 200):
        xb = O.make_activations(M, K, seed=M)
        y = bits_of(fused.forward(t_bf16_from_bits(xb).to(DEV)))
        n0 = 0
        for (_, (qweight, qzeros, scales, b)), n in zip(cks, parts):
            q, z, sb = awq_unpacked(qweight, qzeros, scales)
            want = oracle_forward(xb, q, z, sb, g, bias_f32=O.np_bf16_round(b))
            assert_close_ref(y[:, n0:n0 + n], want, f"part at {n0} M={M}")
            n0 += n


@pytest.mark.parametrize("act_order", [False, True])
def test_gptq_parts_equal_concatenated_checkpoint(product, act_order):
    K, parts, g = 1024, (256, 128), 128
    g_idx = None
    if act_order:
        g_idx = np.random.default_rng(5).permutation(np.repeat(np.arange(K // g, dtype=np.int32), g)).astype(np.int32)
    ws = [gptq_weights(K, n, g, seed=40 + i, g_idx=g_idx)[0] for i, n in enumerate(parts)]
    cat = {"qweight": torch.cat([w["qweight"] for w in ws], dim=1), "scales": torch.cat([w["scales"] for w in ws], dim=1),
           "qzeros": torch.cat([w["qzeros"] for w in ws], dim=1)}
    if act_order:
        cat["g_idx"] = ws[0]["g_idx"]
    cfg = product.GptqConfig(4, g, act_order)
    whole = cfg.create_linear(K, sum(parts), False, DEV)
    whole.load_weights(cat)
    fused = cfg.create_linear(K, sum(parts), False, DEV)
    fused.load_part(0, ws[0])
    fused.load_part(parts[0], ws[1])
    same_handle(fused, whole)
    for M in (1, 5, 64, 257):
        x = t_bf16_from_bits(O.make_activations(M, K, seed=M)).to(DEV)
        assert (bits_of(fused.forward(x)) == bits_of(whole.forward(x))).all()
    if act_order:   # siblings share x: a different g_idx on one part is refused
        other = dict(ws[1]); other["g_idx"] = torch.from_numpy(np.roll(g_idx, 1))
        bad = cfg.create_linear(K, sum(parts), False, DEV)
        bad.load_part(0, ws[0])
        with pytest.raises(product.W4A16Error, match="same g_idx"):
            bad.load_part(parts[0], other)


def test_part_load_errors(product):
    K, g = 512, 128
    cfg = product.AwqConfig.int4(g)
    w64 = awq_weights(K, 64, g, 1)[0]
    w32 = awq_weights(K, 32, g, 2)[0]
    lin = cfg.create_linear(K, 128, False, DEV)
    lin.load_part(0, w64)
    x = torch.zeros((1, K), dtype=torch.bfloat16, device=DEV)
    with pytest.raises(product.W4A16Error, match="not loaded"):     # half the columns are still empty
        lin.forward(x)
    with pytest.raises(product.W4A16Error, match="overlaps"):
        lin.load_part(32, w64)
    with pytest.raises(product.W4A16Error, match="out of range"):
        lin.load_part(96, w64)
    with pytest.raises(product.W4A16Error, match="multiples of 16"):
        lin.load_part(72, w32)
    lin.load_part(0, w64)            # the same part again: overwritten in place
    lin.load_part(64, w64)
    lin.forward(x)
    with pytest.raises(product.W4A16Error, match="before any weights"):
        lin.set_epilogue(product.EPILOGUE_SWIGLU)
    sw = cfg.create_linear(K, 128, False, DEV)
    sw.set_epilogue(product.EPILOGUE_SWIGLU)
    with pytest.raises(product.W4A16Error, match="load_\\*_part|_part"):
        sw.load_weights(awq_weights(K, 128, g, 3)[0])
    with pytest.raises(product.W4A16Error, match="two parts"):
        sw.load_part(0, w32)
    with pytest.raises(product.W4A16Error, match="two parts"):
        sw.load_part(32, w64)
    shard = cfg.create_linear(K, 128, False, DEV, tp_mode=1, tp_rank=0, tp_size=2)
    with pytest.raises(product.W4A16Error, match="unsharded"):
        shard.load_part(0, w32)
    with pytest.raises(product.W4A16Error, match="unsharded"):
        shard.set_epilogue(product.EPILOGUE_SWIGLU)


# ---- SwiGLU epilogue ---------------------------------------------------------------------------------------------------
def exact_awq(K, N, g, seed):
    """An AWQ checkpoint with power-of-two scales: with small-integer activations every product and every partial sum
    is exactly representable in fp32, so the accumulated projection is the same number in ANY summation order."""
    rng = np.random.default_rng(seed)
    qweight, qzeros, _, _ = O.make_awq_checkpoint(K, N, g, seed=seed)
    scales = (2.0 ** rng.integers(-9, -6, size=(K // g, N))).astype(np.float16)
    return {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}


def exact_x(M, K, seed):
    rng = np.random.default_rng(seed)
    return torch.from_numpy(rng.integers(-2, 3, size=(M, K)).astype(np.float32)).to(torch.bfloat16)


def swiglu_layer(product, K, I, g, wg, wu, bias=False):
    lin = product.AwqConfig.int4(g).create_linear(K, 2 * I, bias, DEV)
    lin.set_epilogue(product.EPILOGUE_SWIGLU)
    lin.load_part(I, wu)       # up first: the order of the two loads is free
    lin.load_part(0, wg)
    assert lin.out_features() == I and lin.weight_rows() == 2 * I
    return lin


@pytest.mark.parametrize("K,I,g", [(4096, 1536, 128),    # K = 4096: gemv_imma2 / decode / prefill
                                   (2048, 768, 64),      # gemv_imma v1
                                   (1024, 2048 + 64, 32),  # I not a multiple of 64 x 2: half-full last prefill tile
                                   (512, 64, 128)])      # tiny: the stream kernel
def test_swiglu_bit_exact_against_the_reference_kernels(product, K, I, g):
    ref = _ref()
    wg, wu = exact_awq(K, I, g, 11), exact_awq(K, I, g, 12)
    cfg = product.AwqConfig.int4(g)
    gate, up = cfg.create_linear(K, I, False, DEV), cfg.create_linear(K, I, False, DEV)
    gate.load_weights(wg); up.load_weights(wu)
    fused = swiglu_layer(product, K, I, g, wg, wu)
    # the hooks of a SwiGLU handle answer in [gate | up] column order
    q, z = fused.unpack()
    qg, zg = gate.unpack(); qu, zu = up.unpack()
    assert torch.equal(q, torch.cat([qg, qu], 1)) and torch.equal(z, torch.cat([zg, zu], 1))
    assert (bits_of(fused.scales()) == bits_of(torch.cat([gate.scales(), up.scales()], 1))).all()
    assert (bits_of(fused.dequantize()) == bits_of(torch.cat([gate.dequantize(), up.dequantize()], 1))).all()
    kernels = set()
    for M in M_LIST:
        x = exact_x(M, K, seed=M).to(DEV)
        a, b = gate.forward(x), up.forward(x)
        y = torch.full((M, I), float("nan"), dtype=torch.bfloat16, device=DEV)
        fused.forward_into(x, y)
        torch.cuda.synchronize()
        kernels.add(fused.kernel_for(M))
        want_np = O.np_swiglu_bf16(bits_of(a), bits_of(b))
        got = bits_of(y)
        if ref is not None:
            s = torch.cuda.current_stream().cuda_stream
            for fn in (ref.ref_silu_and_mul_separate_bf16, ref.ref_fused_swiglu_bf16):
                want = torch.empty_like(y)
                assert fn(want.data_ptr(), a.data_ptr(), b.data_ptr(), M, I, s) == 0
                torch.cuda.synchronize()
                assert (got == bits_of(want)).all(), f"M={M} kernel {fused.kernel_for(M)}: differs from the reference kernel"
        # exact-libm oracle: at most one bf16 ulp apart, on few elements
        d = np.abs(got.astype(np.int32) - want_np.astype(np.int32))
        assert d.max() <= 1 and (d != 0).mean() < 0.02, f"M={M}: {d.max()} ulp, {(d != 0).mean():.4f} of elements"
    assert len(kernels) >= 2, kernels


@pytest.mark.parametrize("bias", [False, True])
def test_swiglu_random_inputs_within_tolerance(product, bias):
    K, I, g = 4096, 1024, 128
    (wg, ckg), (wu, cku) = awq_weights(K, I, g, 21, bias), awq_weights(K, I, g, 22, bias)
    fused = swiglu_layer(product, K, I, g, wg, wu, bias)
    for M in (1, 3, 8, 32, 256):
        xb = O.make_activations(M, K, seed=M)
        y = fused.forward(t_bf16_from_bits(xb).to(DEV))
        proj = []
        for qweight, qzeros, scales, b in (ckg, cku):
            q, z, sb = awq_unpacked(qweight, qzeros, scales)
            want = oracle_forward(xb, q, z, sb, g, bias_f32=None if b is None else O.np_bf16_round(b))
            proj.append(O.np_f32_to_bf16_bits(want.astype(np.float32)))
        want = O.np_bf16_bits_to_f32(O.np_swiglu_bf16(proj[0], proj[1])).astype(np.float64)
        got = O.np_bf16_bits_to_f32(bits_of(y)).astype(np.float64)
        # two factors, each within the linear's bound (1e-2 of its scale), then one more bf16 rounding: the product is
        # held to 2e-2 of its own scale
        err = np.abs(got - want).max()
        assert err <= 2e-2 * np.abs(want).max(), f"M={M}: {err} vs max {np.abs(want).max()}"


def test_swiglu_through_the_loader_matches_separate_linears(product):
    """load_swiglu / load_fused on a VarBuilder-style mapping, as QuantizedSwiGluMlp::new pulls gate_proj / up_proj
    (qwen3_quantized.rs:501-520)."""
    K, I, g = 1024, 512, 128
    wg, wu = exact_awq(K, I, g, 31), exact_awq(K, I, g, 32)
    tensors = {f"mlp.gate_proj.{k}": v for k, v in wg.items()} | {f"mlp.up_proj.{k}": v for k, v in wu.items()}
    det = product.DetectedQuantConfig(product.QuantizationMethod.AWQ, 4, g, None, {})
    loader = product.create_weight_loader_with_params(tensors, det, DEV)
    fused = loader.load_swiglu("mlp.gate_proj", "mlp.up_proj", K, I)
    both = loader.load_fused(["mlp.gate_proj", "mlp.up_proj"], K, [I, I], False)
    gate, up = loader.load_linear("mlp.gate_proj", K, I, False), loader.load_linear("mlp.up_proj", K, I, False)
    for M in (1, 6, 100):
        x = exact_x(M, K, seed=M).to(DEV)
        a, b = gate.forward(x), up.forward(x)
        assert (bits_of(both.forward(x)) == bits_of(torch.cat([a, b], 1))).all()
        want = O.np_swiglu_bf16(bits_of(a), bits_of(b))
        d = np.abs(bits_of(fused.forward(x)).astype(np.int32) - want.astype(np.int32))
        assert d.max() <= 1
