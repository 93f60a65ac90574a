"""Pins the CPU oracle to the reference's own known-answer tests.

Each test names the reference test it restates (file:line under
/root/reference/crates/core/src/quantization/).  CPU only.
"""
import numpy as np
import pytest

from oracle import oracle as O

C = O.C
AWQ_PACK_SHIFTS = [0, 16, 4, 20, 8, 24, 12, 28]  # awq_marlin.rs:372


def awq_pack_from_fn(K, N, fn):
    awq = np.zeros((K, N // 8), np.uint32)
    for k in range(K):
        for n in range(N):
            awq[k, n // 8] |= np.uint32(fn(k, n) << AWQ_PACK_SHIFTS[n % 8])
    return awq


# ---- K-1: awq_marlin.rs:721-738 ------------------------------------------------
def test_k1_awq_to_gptq_u32_known_values():
    assert C.awq_to_gptq_u32(0) == 0
    awq = 1 | (3 << 4) | (5 << 8) | (7 << 12) | (2 << 16) | (4 << 20) | (6 << 24) | (8 << 28)
    assert C.awq_to_gptq_u32(awq) == 0x87654321


def test_k1_repack_awq_nibbles_shape_and_zero():
    # awq_marlin.rs:741-748, 801-809
    z = np.zeros((16, 64), np.uint32)
    out = C.repack_awq_nibbles(z)
    assert out.shape == (16, 64) and not out.any()


# ---- K-2: awq_marlin.rs:750-799 ------------------------------------------------
def test_k2_awq_to_gptq_qweight_known_value():
    k = n = 8
    words = np.zeros((k, 1), np.uint32)
    for ki in range(k):
        v = (ki + 1) & 0xF
        words[ki, 0] = sum(v << s for s in range(0, 32, 4))
    out = C.awq_to_gptq_qweight(words, k, n)
    assert out.shape == (1, 8)
    expected = sum(((ki + 1) & 0xF) << (ki * 4) for ki in range(k))
    assert expected == 0x87654321
    assert (out == expected).all()


def test_k2_shape_transposes_packing_axis():
    out = C.awq_to_gptq_qweight(np.zeros((64, 4), np.uint32), 64, 32)
    assert out.shape == (8, 32)
    with pytest.raises(ValueError):
        C.awq_to_gptq_qweight(np.zeros((7, 1), np.uint32), 7, 8)


# ---- K-3: awq_marlin.rs:875-1000 -----------------------------------------------
@pytest.mark.parametrize("K,N,fn", [
    (16, 64, lambda r, c: ((r * 64 + c) * 7 + 3) % 16),
    (32, 128, lambda r, c: (r * 1023 + c * 17 + 5) % 16),
])
def test_k3_marlin_tile_repack_round_trip(K, N, fn):
    awq = awq_pack_from_fn(K, N, fn)
    out = C.awq_to_marlin_tile_repack(awq, K, N)
    assert out.shape == (K // 16, N * 2)
    for k in range(K):
        for n in range(N):
            idx, bit = C.marlin_tile_decode_position(k % 16, n % 64)
            raw = int(out[k // 16, (n // 64) * 128 + idx])
            assert (raw >> bit) & 0xF == fn(k, n), (k, n)


def test_k3_marlin_tile_repack_shape_and_rejects():
    # awq_marlin.rs:889-917
    out = C.awq_to_marlin_tile_repack(np.zeros((32, 16), np.uint32), 32, 128)
    assert out.shape == (2, 256) and not out.any()
    with pytest.raises(ValueError):
        C.awq_to_marlin_tile_repack(np.zeros((15, 8), np.uint32), 15, 64)
    with pytest.raises(ValueError):
        C.awq_to_marlin_tile_repack(np.zeros((16, 5), np.uint32), 16, 40)


# ---- K-4: awq_marlin.rs:1005-1044 (LCG words; CPU side of the GPU==CPU test) ----
def test_k4_lcg_and_repack_self_consistency():
    k, n = 64, 128
    words = C.lcg_words(0x9E3779B9, k * (n // 8))
    # first LCG output, by hand: 0x9E3779B9*1664525+1013904223 mod 2^32
    assert int(words[0]) == (0x9E3779B9 * 1664525 + 1013904223) & 0xFFFFFFFF
    awq = words.reshape(k, n // 8)
    gptq = C.awq_to_gptq_qweight(awq, k, n)
    # The repack is a pure re-indexing: unpacking either layout gives the same nibbles.
    assert (C.gptq_unpack_qweight(gptq, k, n) == C.awq_unpack(awq, n)).all()
    assert (O.np_gptq_unpack_qweight(gptq) == O.np_awq_unpack(awq)).all()
    # R1 on qzeros-shaped data is the same statement along N.
    assert (O.np_seq_unpack(C.repack_awq_nibbles(awq)) == O.np_awq_unpack(awq)).all()


# ---- K-5: marlin_cuda.rs:1941-2024 (closed form 28.0) ---------------------------
def test_k5_gemv_smoke_closed_form():
    k, n, g = 8, 32, 8
    w_word = sum(i << (i * 4) for i in range(8))
    assert w_word == 0x76543210
    qweight = np.full((k // 8, n), w_word, np.uint32)
    qzeros = np.zeros((k // g, n // 8), np.uint32)
    q = C.gptq_unpack_qweight(qweight, k, n)
    z = C.seq_unpack(qzeros, n)
    y = C.gemm_f32(np.ones((1, k), np.float32), q, z, np.ones((k // g, n), np.float32), None, None, g, k, n)
    assert (y == 28.0).all()


# ---- K-6: marlin_cuda.rs:2032-2612 (LCG cases; oracle self-checks) --------------
K6_CASES = [  # (seed, M, K, N)
    (0x9E3779B9, 1, 256, 128),
    (0xCAFEBABE, 4, 256, 64),
    (0xCAFEBABE ^ 8, 8, 256, 64),
    (0xCAFEBABE ^ 12, 12, 256, 64),
    (0xCAFEBABE ^ 16, 16, 256, 64),
    (0xDEADBEEF, 32, 256, 64),
]


@pytest.mark.parametrize("seed,M,K,N", K6_CASES)
def test_k6_lcg_case_oracle_consistency(seed, M, K, N):
    g = 128
    qw, qz, sc, x = C.lcg_gemm_case(seed, M, K, N, g)
    # generator sanity: bf16-rounded, in the documented ranges
    assert (O.np_bf16_round(sc) == sc).all() and (O.np_bf16_round(x) == x).all()
    assert np.abs(sc).max() <= 0.1 * (1 + 2**-8) and np.abs(x).max() <= 1.0 * (1 + 2**-8)  # bf16 rounding of the range ends
    q, z = C.gptq_unpack_qweight(qw, K, N), C.seq_unpack(qz, N)
    y_c = C.gemm_f32(x, q, z, sc, None, None, g, K, N)
    # pure-Python transcription of marlin_cuda.rs:2091-2112 on a few entries
    rng = np.random.default_rng(0)
    for _ in range(8):
        mi, ni = int(rng.integers(M)), int(rng.integers(N))
        acc = np.float32(0)
        for k in range(K):
            nib = (int(qw[k // 8, ni]) >> ((k % 8) * 4)) & 0xF
            zp = (int(qz[k // g, ni // 8]) >> ((ni % 8) * 4)) & 0xF
            w_val = np.float32(np.float32(nib - zp) * sc[k // g, ni])
            acc = np.float32(acc + np.float32(x[mi, k] * w_val))
        assert y_c[mi, ni] == acc
    # float64 mirror agrees far inside the reference tolerance
    y_np = O.np_gemm_ref(x, q, z, sc, g)
    assert np.abs(y_np - y_c).max() < 1e-3 < O.reference_tolerance(y_c)


# ---- K-8: marlin_tile_cuda.rs:555-584 (nibbles at bits 0,16,4,20) ---------------
def test_k8_lo4_positions():
    qv = (np.arange(256, dtype=np.uint64) * 0x9E3779B1 & 0xFFFFFFFF).astype(np.uint32)
    un = O.np_awq_unpack(qv.reshape(-1, 1))
    # AWQ lanes 0..3 live at bits 0,16,4,20 -- exactly the primitive's outputs
    for j, sh in enumerate([0, 16, 4, 20]):
        assert (un[:, j] == ((qv >> sh) & 0xF)).all()


# ---- K-9: marlin.rs:308-341, 874-919; awq.rs:291-304 (shape gate) ---------------
def test_k9_marlin_shape_gate(product):
    ok = product.check_marlin_supports_shape
    ok(4096, 4096, 128)
    for n, k in [(100, 4096), (4096, 100)]:
        with pytest.raises(ValueError):
            ok(n, k, 128)
    with pytest.raises(ValueError):
        ok(4096, 4096, 96)  # group does not divide K (marlin.rs:332-338)
    assert product.MARLIN_SUPPORTED_GROUP_SIZES == (-1, 32, 64, 128)


# ---- D1 / D2 restatements: numpy mirror == C, f16/bf16 casts == numpy's --------
def test_f16_bf16_casts_match_numpy():
    rng = np.random.default_rng(1)
    v = np.concatenate([rng.normal(0, 1, 4000), rng.normal(0, 1e-6, 2000), rng.normal(0, 3e4, 2000),
                        [0.0, -0.0, 65504.0, 65519.9, 65520.0, 1e-8, 5.96e-8, 2.98e-8, 2.9802322387695312e-08]]
                       ).astype(np.float32)
    lib = C.lib
    got16 = np.array([lib.oracle_f32_to_f16(float(x)) for x in v], np.uint16)
    with np.errstate(over="ignore"):
        assert (got16 == v.astype(np.float16).view(np.uint16)).all()
    back = np.array([lib.oracle_f16_to_f32(int(h)) for h in got16], np.float32)
    assert (back.view(np.uint32) == got16.view(np.float16).astype(np.float32).view(np.uint32)).all()
    gotb = np.array([lib.oracle_f32_to_bf16(float(x)) for x in v], np.uint16)
    assert (gotb == O.np_f32_to_bf16_bits(v)).all()
    import torch
    tb = torch.from_numpy(v).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert (gotb == tb).all()


@pytest.mark.parametrize("K,N,g", [(256, 128, 128), (128, 64, 32), (256, 64, 64)])
def test_d1_d2_numpy_equals_c(K, N, g):
    qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=7)
    q, z = C.awq_unpack(qweight, N), C.awq_unpack(qzeros, N)
    assert (q == O.np_awq_unpack(qweight)).all() and (z == O.np_awq_unpack(qzeros)).all()
    assert (O.np_awq_pack(q) == qweight).all()
    # D1: awq_dequantize_cpu -> f16
    d1 = C.awq_dequantize_cpu(qweight, scales.astype(np.float32), qzeros, g, K, N)
    gi = np.arange(K) // g
    want = ((q.astype(np.int32) - z[gi].astype(np.int32)).astype(np.float32) * scales.astype(np.float32)[gi]).astype(np.float16)
    assert (d1.view(np.uint16) == want.view(np.uint16)).all()
    # D2: bf16 scales (S1), bf16_rn(f32(q-z) * f32(s))
    sb = np.empty(scales.shape, np.uint16)
    C.lib.oracle_f16_to_bf16_array(scales.view(np.uint16).ctypes.data, sb.ctypes.data, sb.size)
    assert (sb == O.np_f32_to_bf16_bits(scales.astype(np.float32))).all()
    d2_c = C.dequant_bf16(q, z, sb, None, g, K, N)
    assert (d2_c == O.np_dequant_bf16(q, z, sb, g)).all()
    # symmetric (G3) and act-order (G4) variants
    assert (C.dequant_bf16(q, z, sb, None, g, K, N, 1) == O.np_dequant_bf16(q, z, sb, g, zero_mode=1)).all()
    g_idx = np.random.default_rng(3).permutation(np.repeat(np.arange(K // g, dtype=np.int32), g)).astype(np.int32)
    assert (C.dequant_bf16(q, z, sb, g_idx, g, K, N) == O.np_dequant_bf16(q, z, sb, g, g_idx=g_idx)).all()
    triv = (np.arange(K) // g).astype(np.int32)
    assert (C.dequant_bf16(q, z, sb, triv, g, K, N) == d2_c).all()  # the pinned sub-case of G4


def test_d2_equals_single_rounded_bf16_product_exhaustive():
    """DESIGN claim used by the kernels: bf16_rn(f32(q-z)*f32(s)) == the correctly
    rounded bf16 product of the exact integer (q-z) and the bf16 scale, for every
    (q-z) in [-15,15] and every finite bf16 scale.  torch's bf16 mul is that product."""
    import torch
    s_bits = np.arange(65536, dtype=np.uint32).astype(np.uint16)
    s = O.np_bf16_bits_to_f32(s_bits)
    fin = np.isfinite(s)
    st = torch.from_numpy(s_bits[fin].view(np.int16)).view(torch.bfloat16)
    for d in range(-15, 16):
        want = O.np_f32_to_bf16_bits(np.float32(d) * s[fin])
        got = (torch.tensor(float(d), dtype=torch.bfloat16) * st).view(torch.int16).numpy().view(np.uint16)
        assert (want == got).all(), d


def test_f1_awq_linear_forward_threads_agree_and_matches_f64():
    K, N, g, M = 256, 128, 128, 3
    qweight, qzeros, scales, bias = O.make_awq_checkpoint(K, N, g, seed=11, bias=True)
    x = O.make_activations(M, K, seed=5)
    bb = O.np_f32_to_bf16_bits(bias)
    sf = scales.astype(np.float32)
    y1 = C.awq_linear_forward(x, qweight, sf, qzeros, bb, g, K, N, threads=1)
    y4 = C.awq_linear_forward(x, qweight, sf, qzeros, bb, g, K, N, threads=4)
    assert (y1 == y4).all()
    w = C.awq_dequantize_cpu(qweight, sf, qzeros, g, K, N).astype(np.float64)
    ref = O.np_bf16_bits_to_f32(x).astype(np.float64) @ w + O.np_bf16_bits_to_f32(bb).astype(np.float64)
    got = O.np_bf16_bits_to_f32(y1)
    assert np.abs(got - ref).max() <= O.reference_tolerance(ref)
