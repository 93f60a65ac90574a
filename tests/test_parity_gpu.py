"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle.

Integer work (unpack, repack round trip, bf16 scales) and the dequantised weights
are checked BIT-EXACT.  The GEMM is checked with the reference's own tolerance
max(0.01*max|y|, 5e-2) (marlin_cuda.rs:2168-2183) and the north-star bound
max|err| <= 1e-2*max|y| against the fp32 dequant-matmul on bf16-rounded operands."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import (assert_close_ref, awq_unpacked, bits_of, gptq_unpacked, oracle_forward, t_bf16_from_bits, t_i32)

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def make_awq_layer(P, K, N, g, seed, bias=False, **tp):
    qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g if g > 0 else K, seed=seed, bias=bias)
    cfg = P.AwqConfig.int4(g)
    lin = cfg.create_linear(K, N, bias, DEV, **tp)
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
    if bias:
        w["bias"] = torch.from_numpy(b).to(torch.bfloat16)
    lin.load_weights(w)
    bias_r = None if not bias else O.np_bf16_round(b)
    return lin, (qweight, qzeros, scales, bias_r)


def run_forward(lin, x_bits):
    x = t_bf16_from_bits(x_bits).to(DEV)
    y = lin.forward(x)
    torch.cuda.synchronize()
    return bits_of(y)


# ---- integer / dequant: bit exact ---------------------------------------------------
@pytest.mark.parametrize("K,N,g", [(128, 64, 128), (256, 128, 128), (512, 192, 64), (256, 64, 32), (384, 128, -1),
                                   (1024, 6144, 128)])
def test_awq_unpack_scales_dequant_bit_exact(product, K, N, g):
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=K + N)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    gq, gz = lin.unpack()
    assert (gq.cpu().numpy() == q).all(), "qweight nibbles differ"
    assert (gz.cpu().numpy() == z).all(), "zero points differ"
    assert (bits_of(lin.scales()) == sb).all(), "bf16 scales differ"
    want = O.np_dequant_bf16(q, z, sb, g if g > 0 else K)
    assert (bits_of(lin.dequantize()) == want).all(), "dequantised weights differ (D2 grid)"
    if K * N <= 256 * 128:  # and against the scalar C restatement of awq_marlin_dequant.cu
        assert (want == O.C.dequant_bf16(q, z, sb, None, g if g > 0 else K, K, N)).all()


@pytest.mark.parametrize("sym", [False, True])
@pytest.mark.parametrize("act_order", [False, True])
def test_gptq_unpack_dequant_bit_exact(product, sym, act_order):
    K, N, g = 512, 128, 128
    qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, g, seed=3, act_order=act_order)
    cfg = product.GptqConfig(4, g, desc_act=act_order, use_stored_zeros=not sym)
    lin = cfg.create_linear(K, N, False, DEV)
    w = {"qweight": t_i32(qweight), "scales": torch.from_numpy(scales), "qzeros": t_i32(qzeros)}
    if act_order:
        w["g_idx"] = torch.from_numpy(g_idx)
    lin.load_weights(w)
    q, z, sb = gptq_unpacked(qweight, qzeros, scales)
    gq, gz = lin.unpack()
    assert (gq.cpu().numpy() == q).all()
    if not sym:
        assert (gz.cpu().numpy() == z).all()
    else:
        assert (gz.cpu().numpy() == 8).all()
    want = O.np_dequant_bf16(q, z, sb, g, g_idx=g_idx, zero_mode=1 if sym else 0)
    assert (bits_of(lin.dequantize()) == want).all()


# ---- reference KATs through the CUDA path ----------------------------------------------
def test_k5_gemv_smoke_known_values(product):
    """marlin_cuda.rs:1941-2024 restated at the smallest shape the gate admits: every word
    0x76543210 along K, z = 0, s = 1, x = 1 -> each group of 8 k sums to 28."""
    K, N, g = 128, 64, 128
    qweight = np.full((K // 8, N), 0x76543210, np.uint32)
    qzeros = np.zeros((1, N // 8), np.uint32)
    lin = product.GptqConfig(4, g).create_linear(K, N, False, DEV)
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.ones((1, N), dtype=torch.float16)})
    y = lin.forward(torch.ones((1, K), dtype=torch.bfloat16, device=DEV))
    assert (y.float().cpu().numpy() == 28.0 * (K // 8)).all()


K6_CASES = [(0x9E3779B9, 1, 256, 128), (0xCAFEBABE, 4, 256, 64), (0xCAFEBABE ^ 8, 8, 256, 64),
            (0xCAFEBABE ^ 12, 12, 256, 64), (0xCAFEBABE ^ 16, 16, 256, 64), (0xDEADBEEF, 32, 256, 64)]


@pytest.mark.parametrize("seed,M,K,N", K6_CASES)
def test_k6_lcg_cases_match_scalar_oracle(product, seed, M, K, N):
    """test_awq_gemv_matches_cpu_reference_m{1,4,8,12,16} / ..._m32 (marlin_cuda.rs:2032-2612):
    same LCG data, same scalar fp32 oracle, same tolerance."""
    g = 128
    qw, qz, sc, x = O.C.lcg_gemm_case(seed, M, K, N, g)
    lin = product.GptqConfig(4, g).create_linear(K, N, False, DEV)
    lin.load_weights({"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc)})
    q, z = O.C.gptq_unpack_qweight(qw, K, N), O.C.seq_unpack(qz, N)
    want = O.C.gemm_f32(x, q, z, sc, None, None, g, K, N)
    got = run_forward(lin, O.np_f32_to_bf16_bits(x))
    err = np.abs(O.np_bf16_bits_to_f32(got) - want).max()
    assert err <= O.reference_tolerance(want), err
    assert err <= 1e-2 * np.abs(want).max() + 1e-6


def test_k7_tile_mma_lockdown_data(product):
    """marlin_tile_cuda.rs:594-726 data: a=((i*0.137)%2)-1, nib=(113k+29n+7)%16, zp=(17g+11n+3)%16,
    s=0.05+0.01g+0.0025n, tightest tolerance of the nine shapes (5e-2 abs is for K=16; scaled by max|y|)."""
    for M, K, N in [(4, 128, 64), (16, 256, 128), (32, 512, 64)]:
        g = 128
        k, n = np.arange(K)[:, None], np.arange(N)[None, :]
        q = ((113 * k + 29 * n + 7) % 16).astype(np.uint8)
        gi = np.arange(K // g)[:, None]
        z = ((17 * gi + 11 * n + 3) % 16).astype(np.uint8)
        s = (0.05 + 0.01 * gi + 0.0025 * n).astype(np.float32)
        a = (((np.arange(M * K) * 0.137) % 2) - 1).astype(np.float32).reshape(M, K)
        xb = O.np_f32_to_bf16_bits(a)
        lin = product.GptqConfig(4, g).create_linear(K, N, False, DEV)
        lin.load_weights({"qweight": t_i32(O.np_gptq_pack_qweight(q)), "qzeros": t_i32(O.np_seq_pack(z)),
                          "scales": torch.from_numpy(s)})
        sb = O.np_f32_to_bf16_bits(s)
        want = oracle_forward(xb, q, z, sb, g)
        assert_close_ref(run_forward(lin, xb), want, f"K-7 {M}x{K}x{N}")


# ---- GEMM parity across M, group sizes, bias ----------------------------------------------
@pytest.mark.parametrize("M", [1, 2, 3, 7, 8, 9, 15, 16, 17, 31, 32, 33, 64, 100])
def test_awq_forward_all_m(product, M):
    K, N, g = 1024, 256, 128
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=17)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=M)
    assert_close_ref(run_forward(lin, xb), oracle_forward(xb, q, z, sb, g), f"M={M}")


@pytest.mark.parametrize("g", [32, 64, 128, -1])
@pytest.mark.parametrize("M", [1, 16, 32])
def test_awq_forward_group_sizes_and_bias(product, g, M):
    K, N = 512, 192
    lin, (qweight, qzeros, scales, bias) = make_awq_layer(product, K, N, g, seed=5 + M, bias=True)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=11)
    assert_close_ref(run_forward(lin, xb), oracle_forward(xb, q, z, sb, g if g > 0 else K, bias_f32=bias), f"g={g}")


@pytest.mark.parametrize("dispatch", ["decode", "prefill"])
@pytest.mark.parametrize("M", [5, 17, 32, 64, 128])
def test_forced_dispatch_every_kernel_family(product, dispatch, M):
    """Every kernel family on the same inputs (auto picks one per M; the others must agree with the oracle too).
    K = 2048 gives the split-K / multi-iteration paths of the tcgen05 kernels something to do."""
    K, N, g = 2048, 512, 128
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=23)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=100 + M)
    product.set_dispatch(dispatch)
    try:
        got = run_forward(lin, xb)
    finally:
        product.set_dispatch("auto")
    assert_close_ref(got, oracle_forward(xb, q, z, sb, g), f"{dispatch} M={M}")


@pytest.mark.parametrize("g,K,N,M", [(32, 512, 192, 17), (64, 1024, 320, 32), (-1, 640, 192, 21), (128, 512, 40960, 24),
                                     (128, 12288, 4096, 15), (128, 4096, 6144, 32), (64, 2560, 19456, 29), (128, 4096, 4096, 48), (32, 1024, 2368, 41), (128, 4096, 4096, 11), (64, 2048, 4736, 37)])
def test_midm_kernel_groups_bias_slices_and_many_row_blocks(product, g, K, N, M):
    """The mid-M kernel (11 <= M <= 48, w4a16_midm.cu): group sizes, per-channel scales, bias, K that does not divide
    into equal slices (640 = 5 units, 2560 = 20), N % 128 != 0, more row blocks than warps on the chip (N = 40960:
    several row blocks per warp, the ring runs across them), the Qwen3 shapes; deterministic (slices summed in order)."""
    lin, (qweight, qzeros, scales, bias) = make_awq_layer(product, K, N, g, seed=11 + M, bias=True)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=5)
    assert lin.kernel_for(M) == "midm_mma"
    got = run_forward(lin, xb)
    again = run_forward(lin, xb)
    assert_close_ref(got, oracle_forward(xb, q, z, sb, g if g > 0 else K, bias_f32=bias), f"midm g={g} K={K} N={N} M={M}")
    assert (got == again).all()


def test_midm_kernel_inside_a_graph_capture_without_scratch_falls_back(product):
    """A captured forward on a stream that has no split-K scratch yet must not allocate: it takes the kernels that need
    none, and the replay matches the un-captured result within the GEMM tolerance."""
    K, N, M, g = 4096, 4096, 24, 128
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=41)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=2)
    x = t_bf16_from_bits(xb).to(DEV)
    y = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    st = torch.cuda.Stream(DEV)   # a fresh stream: no (device, stream) scratch exists for it
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st):
        with torch.cuda.graph(graph, stream=st):
            lin.forward_into(x, y)
        graph.replay()
        st.synchronize()
    assert_close_ref(bits_of(y), oracle_forward(xb, q, z, sb, g), "midm capture fallback")


@pytest.mark.parametrize("g,K,N,M", [(32, 512, 192, 57), (64, 1024, 320, 128), (-1, 512, 192, 61), (128, 512, 40960, 50)])
def test_prefill_kernel_groups_bias_tail_and_many_tiles(product, g, K, N, M):
    """The tcgen05 kernel at few tokens: group sizes, per-channel scales, bias, a half-full last 128-feature tile
    (N % 128 != 0), and more tiles than SMs (N = 40960: persistent CTAs run several tiles); deterministic."""
    lin, (qweight, qzeros, scales, bias) = make_awq_layer(product, K, N, g, seed=3 + M, bias=True)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=7)
    assert lin.kernel_for(M) == "prefill_umma"
    got = run_forward(lin, xb)
    again = run_forward(lin, xb)
    assert_close_ref(got, oracle_forward(xb, q, z, sb, g if g > 0 else K, bias_f32=bias), f"prefill g={g} N={N}")
    assert (got == again).all()   # split-K partials are summed in slice order: deterministic


@pytest.mark.parametrize("M", [5, 8])
@pytest.mark.parametrize("K,N,g", [(4096, 4096, 128), (2048, 4736, 64), (1024, 2368, 32)])
def test_imma_two_token_tiles_opt_in(product, M, K, N, g):
    """The int8 tensor-core GEMV with two token tiles (5..8 tokens), opt-in through W4A16_IMMA8 (read per call)."""
    import os
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=41 + M)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=200 + M)
    os.environ["W4A16_IMMA8"] = "1"
    try:
        assert lin.kernel_for(M) == "gemv_stream_imma"
        got = run_forward(lin, xb)
    finally:
        del os.environ["W4A16_IMMA8"]
    assert lin.kernel_for(M) != "gemv_stream_imma"
    assert_close_ref(got, oracle_forward(xb, q, z, sb, g), f"imma8 M={M} {K}x{N} g={g}")


def test_rank3_input_and_noncontiguous(product):
    K, N, g = 256, 128, 128
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=9)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(6, K, seed=2)
    x = t_bf16_from_bits(xb).to(DEV).view(2, 3, K)
    y = lin.forward(x)
    assert tuple(y.shape) == (2, 3, N)
    assert_close_ref(bits_of(y).reshape(6, N), oracle_forward(xb, q, z, sb, g), "rank3")
    xt = t_bf16_from_bits(xb).to(DEV).t().contiguous().t()  # non-contiguous view
    assert not xt.is_contiguous()
    assert (bits_of(lin.forward(xt)) == bits_of(y).reshape(6, N)).all()


def test_determinism_and_every_element_written(product):
    K, N, g = 2048, 128, 128  # short N: the decode kernel splits K across CTAs (ticket path)
    lin, _ = make_awq_layer(product, K, N, g, seed=4)
    x = t_bf16_from_bits(O.make_activations(5, K, seed=3)).to(DEV)
    y0 = torch.full((5, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    lin.forward_into(x, y0)
    assert not torch.isnan(y0.float()).any()
    for _ in range(5):
        y1 = torch.full((5, N), float("nan"), dtype=torch.bfloat16, device=DEV)
        lin.forward_into(x, y1)
        assert (bits_of(y1) == bits_of(y0)).all()


# ---- BASELINE configs at full size (np float64 oracle) ---------------------------------------
@pytest.mark.parametrize("K,N", [(1024, 6144), (4096, 6144), (4096, 4096), (4096, 24576), (12288, 4096)])
def test_config_1_2_shapes_m1(product, K, N):
    g = 128
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=K ^ N)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(1, K, seed=1)
    assert_close_ref(run_forward(lin, xb), oracle_forward(xb, q, z, sb, g), f"{K}x{N}")


_cfg4 = {}


@pytest.mark.parametrize("M", [1, 2, 4, 8, 16, 32, 64, 128, 256])
@pytest.mark.parametrize("K,N", [(2560, 6144), (4096, 2560), (2560, 19456), (9728, 2560)])
def test_config_4_gptq_act_order(product, M, K, N):
    """BASELINE config 4: Qwen3-4B GPTQ int4 with act-order g_idx, all four shapes, the whole M list."""
    g = 128
    if _cfg4.get("key") != (K, N):   # one checkpoint per shape (the shape parameter is the slow axis)
        qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, g, seed=K + N, act_order=True)
        lin = product.GptqConfig(4, g, desc_act=True).create_linear(K, N, False, DEV)
        lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales),
                          "g_idx": torch.from_numpy(g_idx)})
        _cfg4["key"], _cfg4["val"] = (K, N), (lin, gptq_unpacked(qweight, qzeros, scales), g_idx)
    lin, (q, z, sb), g_idx = _cfg4["val"]
    xb = O.make_activations(M, K, seed=M)
    assert_close_ref(run_forward(lin, xb), oracle_forward(xb, q, z, sb, g, g_idx=g_idx), f"act-order {K}x{N} M={M}")


def test_gptq_trivial_g_idx_is_the_pinned_subcase(product):
    K, N, g = 512, 128, 128
    qweight, qzeros, scales, _ = O.make_gptq_checkpoint(K, N, g, seed=1)
    triv = torch.arange(K, dtype=torch.int32) // g
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
    a = product.GptqConfig(4, g).create_linear(K, N, False, DEV)
    a.load_weights(w)
    b = product.GptqConfig(4, g, desc_act=True).create_linear(K, N, False, DEV)
    b.load_weights({**w, "g_idx": triv})
    x = t_bf16_from_bits(O.make_activations(4, K)).to(DEV)
    assert (bits_of(a.forward(x)) == bits_of(b.forward(x))).all()


# ---- TP composition on one GPU (config 5 semantics) ------------------------------------
@pytest.mark.parametrize("tp", [2, 4, 8])
def test_tp_shards_compose(product, tp):
    K, N, g, M = 2048, 1024, 128, 4
    qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=tp, bias=True)
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales),
         "bias": torch.from_numpy(b).to(torch.bfloat16)}
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=8)
    x = t_bf16_from_bits(xb).to(DEV)
    cfg = product.AwqConfig.int4(g)
    # column parallel: concatenated shard outputs == the full layer (the fp32 summation order follows
    # the work partition, which depends on N: equal to within bf16 rounding, each within tolerance)
    full = cfg.create_linear(K, N, True, DEV)
    full.load_weights(w)
    y_full = bits_of(full.forward(x))
    want_full = oracle_forward(xb, q, z, sb, g, bias_f32=O.np_bf16_round(b))
    cols = []
    for r in range(tp):
        sh = cfg.create_linear(K, N, True, DEV, tp_mode=1, tp_rank=r, tp_size=tp)
        sh.load_weights(w)
        assert sh.out_features() == N // tp and sh.in_features() == K
        cols.append(bits_of(sh.forward(x)))
    y_cat = np.concatenate(cols, axis=1)
    assert_close_ref(y_cat, want_full, f"column shards tp={tp}")
    diff = np.abs(O.np_bf16_bits_to_f32(y_cat) - O.np_bf16_bits_to_f32(y_full))
    assert diff.max() <= 2.0 ** -7 * np.abs(want_full).max()
    # row parallel: sum of partials (no bias) + bias == oracle within tolerance
    acc = np.zeros((M, N), np.float64)
    for r in range(tp):
        sh = cfg.create_linear(K, N, True, DEV, tp_mode=2, tp_rank=r, tp_size=tp)
        sh.load_weights(w)
        k0, k1 = product.shard_range(K, r, tp)
        part = bits_of(sh.forward(x[:, k0:k1]))
        want_part = oracle_forward(xb[:, k0:k1], q[k0:k1], z[k0 // g:k1 // g], sb[k0 // g:k1 // g], g)
        assert_close_ref(part, want_part, f"row shard {r}/{tp}")
        acc += O.np_bf16_bits_to_f32(part)
    want = oracle_forward(xb, q, z, sb, g, bias_f32=O.np_bf16_round(b))
    got = acc + O.np_bf16_round(b)[None, :]
    assert np.abs(got - want).max() <= O.reference_tolerance(want)


# ---- error behaviour (marlin_cuda.rs:138-146, 546-564; awq.rs:691-699) ----------------------------
def test_errors(product):
    cfg = product.AwqConfig.int4(128)
    lin = cfg.create_linear(256, 128, False, DEV)
    x = torch.zeros((1, 256), dtype=torch.bfloat16, device=DEV)
    with pytest.raises(product.W4A16Error, match="not loaded"):
        lin.forward(x)
    qweight, qzeros, scales, _ = O.make_awq_checkpoint(256, 128, 128)
    with pytest.raises(product.W4A16Error, match="Missing 'scales'"):
        lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros)})
    with pytest.raises(product.W4A16Error, match="shape"):
        lin.load_weights({"qweight": t_i32(qweight[:, :8]), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    with pytest.raises(product.W4A16Error, match="BF16"):
        lin.forward(x.to(torch.float16))
    with pytest.raises(product.W4A16Error, match="in_features"):
        lin.forward(torch.zeros((1, 128), dtype=torch.bfloat16, device=DEV))
    with pytest.raises(product.W4A16Error):
        cfg.create_linear(256, 100, False, DEV)  # outside the Marlin gate: no CPU fallback


def test_forward_host_matches_forward(product):
    K, N, g, M = 1024, 256, 128, 3
    lin, _ = make_awq_layer(product, K, N, g, seed=2)
    xb = O.make_activations(M, K)
    xh = t_bf16_from_bits(xb).pin_memory()
    yh = torch.empty((M, N), dtype=torch.bfloat16).pin_memory()
    lin.forward_host(xh, yh)
    assert (bits_of(yh) == run_forward(lin, xb)).all()


def test_cuda_graph_capture(product):
    """forward_into never allocates or synchronises: the OutputPool / graph-capture contract."""
    K, N, g, M = 1024, 256, 128, 4
    lin, _ = make_awq_layer(product, K, N, g, seed=6)
    x = t_bf16_from_bits(O.make_activations(M, K)).to(DEV)
    y = torch.zeros((M, N), dtype=torch.bfloat16, device=DEV)
    want = bits_of(lin.forward(x))
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        lin.forward_into(x, y)  # warm-up on the side stream
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            lin.forward_into(x, y)
            lin.forward_into(x, y)
    y.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert (bits_of(y) == want).all()


@pytest.mark.gpu
@pytest.mark.parametrize("M", [17, 64, 100, 256])
@pytest.mark.parametrize("K,N", [(12288, 4096), (12288, 1024), (4096, 4096)])
def test_prefill_split_k_few_tiles(product, M, K, N):
    """Few tokens x few output tiles: the prefill kernel splits K over CTAs (fp32 partials summed in slice order).
    Same oracle and tolerance as every other GEMM case; run twice to check the self-resetting tickets."""
    g = 128
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=K + N + M)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=M)
    want = oracle_forward(xb, q, z, sb, g)
    first = run_forward(lin, xb)
    assert_close_ref(first, want, f"split-K {K}x{N} M={M}")
    assert (run_forward(lin, xb) == first).all(), "split-K result is not deterministic across launches"


# ---- the integer-activation decode kernels (M <= 4): where their arithmetic differs from the reference's ----------
# gemv_imma_kernel rounds activations to 14-bit integers with one power-of-two scale per (group, token)
# (w4a16_gemv_imma.cu); the reference multiplies exact bf16 activations in fp32 (awq_gemv.cu:597-664).
def _imma_case(product, K, N, g, xb, what, seed=77):
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=seed)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    M = xb.shape[0]
    assert lin.kernel_for(M) in ("gemv_stream_imma2", "gemv_stream_imma", "gemv_stream_dp4a", "gemv_stream_nt1"), lin.kernel_for(M)
    return lin, run_forward(lin, xb), oracle_forward(xb, q, z, sb, g)


@pytest.mark.parametrize("M", [1, 4])
@pytest.mark.parametrize("K,N,g", [(4096, 4096, 128), (2048, 4736, 32)])
def test_imma_outlier_per_group(product, M, K, N, g):
    """Every group holds ONE element 2^10 times larger than its 127 (31) small same-sign neighbours: the small ones
    are rounded to ~4 significant bits by the group's scale -- the worst case of the 2^-14 bound in the header."""
    rng = np.random.default_rng(K + M)
    x = np.full((M, K), 0.0123, np.float32) * rng.uniform(0.5, 1.0, size=(M, K)).astype(np.float32)
    for m in range(M):
        pos = rng.integers(0, g, size=K // g) + np.arange(K // g) * g
        x[m, pos] *= 1024.0
    xb = O.np_f32_to_bf16_bits(x)
    _, got, want = _imma_case(product, K, N, g, xb, "outlier")
    assert_close_ref(got, want, f"outlier M={M} g={g}")


@pytest.mark.parametrize("M", [1, 3])
def test_imma_zero_groups_zero_tokens_and_subnormals(product, M):
    K, N, g = 4096, 4096, 128
    xb = O.make_activations(M, K, seed=5)
    xb[:, 256:512] = 0                       # two all-zero groups in every token
    xb[:, 1024:1152] = 0x8000                # a group of negative zeros
    xb[:, 2048:2176] = (np.arange(128) % 127 + 1).astype(np.uint16)[None, :]   # a group of bf16 subnormals
    xb[:, 3000] = 0x0001                     # a subnormal next to normal values
    if M > 1:
        xb[M - 1, :] = 0                     # a whole token of zeros -> exactly zero outputs
    _, got, want = _imma_case(product, K, N, g, xb, "zeros")
    assert_close_ref(got, want, f"zeros/subnormals M={M}")
    if M > 1:
        assert (got[M - 1] & 0x7FFF == 0).all(), "a zero token must give zero outputs"


@pytest.mark.parametrize("M", [1, 4, 8, 64])
def test_non_finite_activations_propagate(product, M):
    """Policy (include/w4a16_b200.h, w4a16_forward_into): a token holding Inf / NaN gets non-finite outputs, as the
    reference's fp32 FMA chain gives it; the other tokens of the batch are untouched.  M = 1, 4: the integer-activation
    kernels (NaN for the whole token); M = 8, 64: the bf16 mma kernels."""
    K, N, g = 4096, 4096, 128
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=31)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=9)
    clean = xb.copy()
    bad = {0: 0x7F80} if M == 1 else {1: 0x7F80, 2: 0x7FC0, M - 1: 0xFF80}   # +Inf, NaN, -Inf
    for m, bits in bad.items():
        xb[m, 777 + m] = bits
    got = O.np_bf16_bits_to_f32(run_forward(lin, xb))
    want_clean = oracle_forward(clean, q, z, sb, g)
    for m in range(M):
        if m in bad:
            assert not np.isfinite(got[m]).any(), f"token {m} holds a non-finite activation: every output must be non-finite"
        else:
            err = np.abs(got[m] - want_clean[m]).max()
            assert err <= 1e-2 * np.abs(want_clean).max(), f"clean token {m} disturbed by a neighbour's Inf/NaN: {err}"


# ---- D1: the reference's CPU dequant grid (awq_dequantize_cpu, awq.rs:56-166) ------------------------------------------
@pytest.mark.parametrize("K,N,g", [(256, 128, 128), (512, 192, 64), (1024, 6144, 128)])
def test_d1_grid_matches_awq_dequantize_cpu_bit_exact(product, K, N, g):
    """f16_rn(f32(q - z) * f32(f16 scale)) with the checkpoint's ORIGINAL scales: a different grid from D2 (bf16-rounded
    scales, bf16 result); SURVEY 8a hazard (i) asks for both to be exposed."""
    lin, (qweight, qzeros, scales, _) = make_awq_layer(product, K, N, g, seed=K + 3 * N)
    want = O.C.awq_dequantize_cpu(qweight, np.asarray(scales, np.float32), qzeros, g, K, N)   # f16 bits [K, N]
    got = lin.dequantize_d1(torch.from_numpy(scales)).cpu().view(torch.int16).numpy().view(np.uint16)
    want_bits = np.asarray(want).view(np.uint16) if np.asarray(want).dtype != np.uint16 else np.asarray(want)
    assert (got == want_bits.reshape(K, N)).all(), "D1 grid differs from awq_dequantize_cpu"
    d2 = O.np_bf16_bits_to_f32(bits_of(lin.dequantize()))
    d1 = got.view(np.float16).astype(np.float32)
    assert np.abs(d1 - d2).max() > 0, "D1 and D2 are different grids (f16 vs bf16 scales): they cannot coincide everywhere"


# ---- act-order + row-parallel: global K permutation before sharding ------------------------------------------
@pytest.mark.parametrize("tp", [2, 4])
def test_act_order_row_shards_compose(product, tp):
    K, N, g, M = 2048, 256, 128, 3
    qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, g, seed=tp + 40, act_order=True)
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales), "g_idx": torch.from_numpy(g_idx)}
    q, z, sb = gptq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=4)
    x = t_bf16_from_bits(xb).to(DEV)
    want = oracle_forward(xb, q, z, sb, g, g_idx=g_idx)
    cfg = product.GptqConfig(4, g, desc_act=True)
    acc = np.zeros((M, N), np.float64)
    seen = []
    for r in range(tp):
        sh = cfg.create_linear(K, N, False, DEV, tp_mode=2, tp_rank=r, tp_size=tp)
        sh.load_weights(w)
        perm = sh.row_perm().numpy()
        assert perm.shape == (K // tp,)
        assert (np.sort(g_idx[perm]) == np.repeat(np.arange(r * K // tp // g, (r + 1) * K // tp // g), g)).all(), \
            "a row shard must hold whole groups of the globally sorted order"
        seen.append(perm)
        part = bits_of(sh.forward(x[:, torch.from_numpy(perm.astype(np.int64)).to(DEV)].contiguous()))
        acc += O.np_bf16_bits_to_f32(part)
    assert (np.sort(np.concatenate(seen)) == np.arange(K)).all(), "the shards' rows must partition K"
    assert np.abs(acc - want).max() <= O.reference_tolerance(want)


@pytest.mark.parametrize("M", [1, 2, 3, 4])
@pytest.mark.parametrize("g", [32, 64, 128])
def test_imma2_k4096_all_group_sizes(product, M, g):
    """The second-generation int8 tensor-core GEMV (K = 4096): hoisted B fragments at g = 128, the shared-memory path
    at g = 32 / 64; bias; a row-block count that is not a multiple of the grid (uneven slots per CTA)."""
    K, N = 4096, 4736
    lin, (qweight, qzeros, scales, bias) = make_awq_layer(product, K, N, g, seed=61 + g + M, bias=True)
    # g = 32 / 64 with 3-4 tokens: the staging no longer fits twice per SM and the plan hands the shape to the first kernel
    assert lin.kernel_for(M) in (("gemv_stream_imma2",) if g == 128 or M <= 2 else ("gemv_stream_imma2", "gemv_stream_imma"))
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=300 + M)
    y = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    lin.forward_into(t_bf16_from_bits(xb).to(DEV), y)
    torch.cuda.synchronize()
    assert_close_ref(bits_of(y), oracle_forward(xb, q, z, sb, g, bias_f32=bias), f"imma2 M={M} g={g}")
