"""a chain of decode linears in one persistent launch (w4a16_chain_*, csrc/w4a16_chain.cu).

The reference runs its quantized linears one launch at a time in a fixed per-layer order (qwen3_quantized.rs:704-709,
532-551) and replays that order as a CUDA graph (engine/cuda_graph.rs).  The chain must give what those separate calls
give.  Checks:
  * every linear of a chain against the float64 oracle (reference tolerance + the 1e-2 north-star bound) and against
    the product's own per-linear w4a16_forward_into (same integer arithmetic, another summation order of the fp32
    partial sums: equal within a few bf16 ulps of max|y|);
  * a TRUE dependency chain (x[i+1] is y[i]): the grid barrier orders the write of one linear before the read of the
    next -- every buffer of the chain must be the linear of the buffer before it, run after run on changing inputs;
  * M = 1..4, group sizes 32 / 64 / 128, bias, the SwiGLU handle (gate/up -> silu*mul -> down as a two-linear chain),
    row blocks fewer than CTAs, repeated runs (the barrier counter is monotonic over runs), CUDA-graph capture;
  * shapes the persistent kernel does not take fall back to per-linear launches with the same results, and say so;
  * argument errors."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import assert_close_ref, awq_unpacked, bits_of, oracle_forward, t_bf16_from_bits, t_i32

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def make_linear(product, K, N, g, seed, bias=False):
    qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=seed, bias=bias)
    lin = product.AwqConfig.int4(g).create_linear(K, N, bias, DEV)
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
    if bias:
        w["bias"] = torch.from_numpy(b).to(torch.bfloat16)
    lin.load_weights(w)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    return lin, (q, z, sb, O.np_bf16_round(b) if bias else None)


@pytest.mark.parametrize("M", [1, 2, 3, 4])
@pytest.mark.parametrize("g", [128, 32])
def test_independent_linears_match_oracle_and_per_linear_calls(product, M, g):
    shapes = [(4096, 1024, False), (8192, 512, True), (4096, 2368 + 64, False), (4096, 256, True)]   # N / 16 < 148 on the last
    lins, cks, xs, ys = [], [], [], []
    for i, (K, N, bias) in enumerate(shapes):
        lin, ck = make_linear(product, K, N, g, seed=10 + i, bias=bias)
        lins.append(lin); cks.append(ck)
        xs.append(t_bf16_from_bits(O.make_activations(M, K, seed=50 + i + M)).to(DEV))
        ys.append(torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV))
    chain = product.DecodeChain(list(zip(lins, xs, ys)))
    assert chain.kernel().startswith("chain_imma") and len(chain) == len(shapes)
    for rep in range(3):                       # the barrier counter keeps counting over runs
        for y in ys:
            y.fill_(float("nan"))
        chain.run()
        torch.cuda.synchronize()
        for i, (K, N, bias) in enumerate(shapes):
            q, z, sb, b = cks[i]
            want = oracle_forward(bits_of(xs[i]), q, z, sb, g, bias_f32=b)
            assert_close_ref(bits_of(ys[i]), want, f"chain linear {i} M={M} g={g} rep={rep}")
            single = lins[i].forward(xs[i]).float()
            assert float((single - ys[i].float()).abs().max()) <= 2.0 ** -6 * float(single.abs().max())


@pytest.mark.parametrize("M", [1, 4])
def test_true_dependency_chain_equals_call_by_call(product, M):
    """x[i+1] = y[i]: qkv-like -> o-like -> gate_up (SwiGLU) -> down, twice."""
    H, I = 4096, 4096
    g = 128
    qweight = {}
    def swiglu(seed):
        lin = product.AwqConfig.int4(g).create_linear(H, 2 * I, False, DEV)
        lin.set_epilogue(product.EPILOGUE_SWIGLU)
        for half in range(2):
            qw, qz, sc, _ = O.make_awq_checkpoint(H, I, g, seed=seed + half)
            # small scales keep the magnitudes of a deep chain in range
            lin.load_part(half * I, {"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc * 0.25)})
        return lin
    def plain(K, N, seed, scale):
        qw, qz, sc, _ = O.make_awq_checkpoint(K, N, g, seed=seed)
        lin = product.AwqConfig.int4(g).create_linear(K, N, False, DEV)
        lin.load_weights({"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc * scale)})
        return lin
    lins = []
    for layer in range(2):
        lins += [plain(H, H, 100 + 10 * layer, 0.3), plain(H, H, 101 + 10 * layer, 0.3), swiglu(102 + 10 * layer),
                 plain(I, H, 105 + 10 * layer, 0.3)]
    x0 = t_bf16_from_bits(O.make_activations(M, H, seed=3)).to(DEV)
    def buffers():
        bufs = [x0.clone()]
        for lin in lins:
            bufs.append(torch.zeros((M, lin.out_features()), dtype=torch.bfloat16, device=DEV))
        return bufs
    a = buffers()
    chain = product.DecodeChain([(lin, a[i], a[i + 1]) for i, lin in enumerate(lins)])
    assert chain.kernel() == "chain_imma (7 mailboxes, 0 barriers)"

    def check(tag):
        torch.cuda.synchronize()
        for i, lin in enumerate(lins):
            want = lin.forward(a[i]).float()          # the linear of what the chain itself left in buffer i
            got = a[i + 1].float()
            assert torch.isfinite(got).all()
            scale = float(want.abs().max())
            assert scale > 0.0
            assert float((want - got).abs().max()) <= 2.0 ** -6 * scale, f"{tag}: buffer {i + 1} is not linear {i} of buffer {i}"

    chain.run()
    check("first run")
    for seed in (4, 5, 6):                      # stale data of the previous run would show
        a[0].copy_(t_bf16_from_bits(O.make_activations(M, H, seed=seed)).to(DEV))
        chain.run()
        check(f"seed {seed}")
    # CUDA-graph capture + replay on fresh inputs
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        chain.run()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            chain.run()
            chain.run()                         # back to back: the second launch may start before the first has ended
        a[0].copy_(t_bf16_from_bits(O.make_activations(M, H, seed=7)).to(DEV))
        graph.replay()
        graph.replay()
    st.synchronize()
    check("graph replay")


def test_mlp_chain_k12288_against_oracle(product):
    """Qwen3-8B MLP shapes: gate_up 4096 -> 2 x 12288 (SwiGLU) then down 12288 -> 4096 (three stages per row block)."""
    H, I, g, M = 4096, 12288, 128, 2
    gu = product.AwqConfig.int4(g).create_linear(H, 2 * I, False, DEV)
    gu.set_epilogue(product.EPILOGUE_SWIGLU)
    for half in range(2):
        qw, qz, sc, _ = O.make_awq_checkpoint(H, I, g, seed=31 + half)
        gu.load_part(half * I, {"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc)})
    down, (q, z, sb, _) = make_linear(product, I, H, g, seed=40)
    x = t_bf16_from_bits(O.make_activations(M, H, seed=9)).to(DEV)
    mid = torch.zeros((M, I), dtype=torch.bfloat16, device=DEV)
    y = torch.zeros((M, H), dtype=torch.bfloat16, device=DEV)
    chain = product.DecodeChain([(gu, x, mid), (down, mid, y)])
    assert chain.kernel() == "chain_imma (1 mailboxes, 0 barriers)"
    chain.run()
    torch.cuda.synchronize()
    mid_ref = gu.forward(x).float()
    assert float((mid_ref - mid.float()).abs().max()) <= 2.0 ** -6 * float(mid_ref.abs().max())
    want = oracle_forward(bits_of(mid), q, z, sb, g)
    assert_close_ref(bits_of(y), want, "down after swiglu")


@pytest.mark.parametrize("M", [1, 3])
def test_ping_pong_buffers_get_barriers(product, M):
    """Two buffers used alternately by linears of different widths: partial overlaps the mailboxes do not cover."""
    g = 128
    shapes = [(4096, 6144), (4096, 4096), (4096, 8192), (8192, 4096), (4096, 6144), (4096, 4096)]
    lins = [make_linear(product, K, N, g, seed=70 + i)[0] for i, (K, N) in enumerate(shapes)]
    width = 8192
    bufs = [torch.zeros((M * width,), dtype=torch.bfloat16, device=DEV) for _ in range(2)]
    def view(b, n):
        return bufs[b][:M * n].view(M, n)
    x0 = t_bf16_from_bits(O.make_activations(M, 4096, seed=11)).to(DEV) * 0.05
    steps = [(lin, view(i & 1, K), view((i + 1) & 1, N)) for i, (lin, (K, N)) in enumerate(zip(lins, shapes))]
    chain = product.DecodeChain(steps)
    name = chain.kernel()
    assert name.startswith("chain_imma") and "0 barriers" not in name
    for rep in range(3):
        view(0, 4096).copy_(x0 * (rep + 1))
        chain.run()
        torch.cuda.synchronize()
        got = view(len(shapes) & 1, shapes[-1][1]).clone()
        # the same sequence call by call (each call reads what the previous one left in the other buffer)
        view(0, 4096).copy_(x0 * (rep + 1))
        for lin, x, y in steps:
            lin.forward_into(x, y)
        torch.cuda.synchronize()
        want = view(len(shapes) & 1, shapes[-1][1]).float()
        scale = float(want.abs().max())
        assert scale > 0 and torch.isfinite(want).all()
        assert float((want - got.float()).abs().max()) <= 0.05 * scale    # six linears of compounding bf16 roundings


@pytest.mark.parametrize("M", [1, 4])
def test_any_k_multiple_of_128(product, M):
    """in_features that are not multiples of 4096 (Qwen3-4B: 2560 / 9728; tensor-parallel shards: 1024, 3584): stages of 32
    units cross row-block boundaries and the last stage of a linear is short."""
    g = 128
    shapes = [(2560, 6144), (1024, 4096), (9728, 2560), (3584, 512), (128, 256), (4096 + 128, 320)]
    lins, cks, xs, ys = [], [], [], []
    for i, (K, N) in enumerate(shapes):
        lin, ck = make_linear(product, K, N, g, seed=60 + i, bias=(i % 2 == 1))
        lins.append(lin); cks.append(ck)
        xs.append(t_bf16_from_bits(O.make_activations(M, K, seed=70 + i)).to(DEV))
        ys.append(torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV))
    chain = product.DecodeChain(list(zip(lins, xs, ys)))
    assert chain.kernel().startswith("chain_imma")
    for rep in range(2):
        chain.run()
        torch.cuda.synchronize()
        for i, (K, N) in enumerate(shapes):
            q, z, sb, b = cks[i]
            assert_close_ref(bits_of(ys[i]), oracle_forward(bits_of(xs[i]), q, z, sb, g, bias_f32=b), f"K={K} N={N} M={M}")


def test_fallback_per_linear(product):
    M = 2
    a, (q, z, sb, _) = make_linear(product, 4096, 512, 64, seed=1)       # group size 64 ...
    b, _ = make_linear(product, 4096, 512, 128, seed=2)                  # ... and 128 in one chain: not taken
    xa = t_bf16_from_bits(O.make_activations(M, 4096, seed=1)).to(DEV)
    xb = t_bf16_from_bits(O.make_activations(M, 4096, seed=2)).to(DEV)
    ya = torch.zeros((M, 512), dtype=torch.bfloat16, device=DEV)
    yb = torch.zeros((M, 512), dtype=torch.bfloat16, device=DEV)
    chain = product.DecodeChain([(a, xa, ya), (b, xb, yb)])
    assert chain.kernel().startswith("per_linear") and "group sizes" in chain.kernel()
    chain.run()
    torch.cuda.synchronize()
    assert (bits_of(ya) == bits_of(a.forward(xa))).all() and (bits_of(yb) == bits_of(b.forward(xb))).all()
    assert_close_ref(bits_of(ya), oracle_forward(bits_of(xa), q, z, sb, 64), "fallback")
    # more than 4 tokens: also per linear
    x8 = t_bf16_from_bits(O.make_activations(8, 4096, seed=3)).to(DEV)
    y8 = torch.zeros((8, 512), dtype=torch.bfloat16, device=DEV)
    c8 = product.DecodeChain([(b, x8, y8)])
    assert c8.kernel().startswith("per_linear")
    c8.run()
    torch.cuda.synchronize()
    assert (bits_of(y8) == bits_of(b.forward(x8))).all()


def test_chain_errors(product):
    lin, _ = make_linear(product, 4096, 256, 128, seed=5)
    x = torch.zeros((1, 4096), dtype=torch.bfloat16, device=DEV)
    y = torch.zeros((1, 256), dtype=torch.bfloat16, device=DEV)
    with pytest.raises(product.W4A16Error, match="at least one"):
        product.DecodeChain([])
    with pytest.raises(product.W4A16Error, match="BF16"):
        product.DecodeChain([(lin, x.float(), y)])
    with pytest.raises(product.W4A16Error, match="in_features"):
        product.DecodeChain([(lin, torch.zeros((1, 2048), dtype=torch.bfloat16, device=DEV), y)])
    with pytest.raises(product.W4A16Error, match="y of step 0"):
        product.DecodeChain([(lin, x, torch.zeros((1, 128), dtype=torch.bfloat16, device=DEV))])
    with pytest.raises(product.W4A16Error, match="same M"):
        product.DecodeChain([(lin, x, y), (lin, torch.zeros((2, 4096), dtype=torch.bfloat16, device=DEV),
                                            torch.zeros((2, 256), dtype=torch.bfloat16, device=DEV))])
    unloaded = product.AwqConfig.int4(128).create_linear(4096, 256, False, DEV)
    with pytest.raises(product.W4A16Error, match="not loaded"):
        product.DecodeChain([(unloaded, x, y)])


# ---- prologues: residual add + RMSNorm inside the chain -------------------------------------------------
import ctypes
import os

REF_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_kernels.so")


def _ref_norm():
    if not os.path.exists(REF_SO):
        return None
    lib = ctypes.CDLL(REF_SO)
    if not hasattr(lib, "ref_rms_norm_bf16") or not hasattr(lib, "ref_add_bf16"):
        return None
    vp, i = ctypes.c_void_p, ctypes.c_int
    lib.ref_rms_norm_bf16.argtypes = [vp, vp, vp, ctypes.c_float, i, i, vp]
    lib.ref_add_bf16.argtypes = [vp, vp, vp, i, i, vp]
    return lib


def _ulp_diff(a_bits, b_bits):
    """distance in bf16 ulps between two bit patterns of the same sign (monotone integer mapping)."""
    def key(b):
        b = b.astype(np.int32)
        return np.where(b & 0x8000, -(b & 0x7FFF), b & 0x7FFF)
    return np.abs(key(a_bits) - key(b_bits))


@pytest.mark.parametrize("M", [1, 2, 4])
def test_prologue_residual_rmsnorm_like_the_reference_layer(product, M):
    """One and a half decoder layers as the reference runs them between attention calls (qwen3_quantized.rs:1043-1075):
         h1 = o_out_prev + residual0 ; gate_up(post_norm(h1)) -> silu*mul -> down ; h2 = down_out + h1 ;
         qkv(input_norm(h2)).
    Chain: [gate_up (prologue: residual_in, norm), down, qkv (prologue: running residual, norm)]."""
    H, I, g, eps = 4096, 4096, 128, 1e-6
    rng = np.random.default_rng(7 + M)
    gu = product.AwqConfig.int4(g).create_linear(H, 2 * I, False, DEV)
    gu.set_epilogue(product.EPILOGUE_SWIGLU)
    gu_ck = []
    for half in range(2):
        qw, qz, sc, _ = O.make_awq_checkpoint(H, I, g, seed=81 + half)
        gu.load_part(half * I, {"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc)})
        gu_ck.append(awq_unpacked(qw, qz, sc))
    down, down_ck = make_linear(product, I, H, g, seed=83)
    qkv, qkv_ck = make_linear(product, H, 6144, g, seed=84)
    attn = t_bf16_from_bits(O.make_activations(M, H, seed=90 + M)).to(DEV)           # o_proj output of the attention block
    res0 = t_bf16_from_bits(O.make_activations(M, H, seed=95 + M)).to(DEV)           # the layer's residual stream
    w_post = torch.from_numpy(rng.uniform(0.5, 1.5, H).astype(np.float32)).to(torch.bfloat16).to(DEV)
    w_in = torch.from_numpy(rng.uniform(0.5, 1.5, H).astype(np.float32)).to(torch.bfloat16).to(DEV)
    z = lambda n: torch.zeros((M, n), dtype=torch.bfloat16, device=DEV)
    mid, dn, y_qkv = z(I), z(H), z(6144)
    h1, n1, h2, n2 = z(H), z(H), z(H), z(H)
    chain = product.DecodeChain(
        [(gu, attn, mid), (down, mid, dn), (qkv, dn, y_qkv)],
        prologues=[{"norm_weight": w_post, "eps": eps, "residual_in": res0, "residual_out": h1, "normed_out": n1},
                   None,
                   {"norm_weight": w_in, "eps": eps, "add_running_residual": True, "residual_out": h2, "normed_out": n2}])
    assert chain.kernel().startswith("chain_imma")
    for rep in range(2):
        chain.run()
        torch.cuda.synchronize()
        # residual adds: elementwise, bit-exact against the restated add_bf16
        assert (bits_of(h1) == O.np_add_bf16(bits_of(attn), bits_of(res0))).all()
        assert (bits_of(h2) == O.np_add_bf16(bits_of(dn), bits_of(h1))).all()
        # norms: against the restated rms_norm_bf16 (float64 sum) within one bf16 ulp, almost everywhere exact
        for hbuf, nbuf, w in ((h1, n1, w_post), (h2, n2, w_in)):
            want = O.np_rms_norm_bf16(bits_of(hbuf), bits_of(w), eps)
            d = _ulp_diff(bits_of(nbuf), want)
            assert d.max() <= 1 and (d == 0).mean() > 0.99, (d.max(), (d == 0).mean())
        # and against the reference's own kernels run on this GPU
        ref = _ref_norm()
        if ref is not None:
            r_h, r_n = z(H), z(H)
            assert ref.ref_add_bf16(r_h.data_ptr(), attn.data_ptr(), res0.data_ptr(), M, H, None) == 0
            assert ref.ref_rms_norm_bf16(r_n.data_ptr(), r_h.data_ptr(), w_post.data_ptr(), eps, M, H, None) == 0
            torch.cuda.synchronize()
            assert (bits_of(r_h) == bits_of(h1)).all()
            d = _ulp_diff(bits_of(n1), bits_of(r_n))
            assert d.max() <= 1 and (d == 0).mean() > 0.99, (d.max(), (d == 0).mean())
        # the linears consume exactly the normalised rows they report
        gate = oracle_forward(bits_of(n1), *gu_ck[0], g)
        up = oracle_forward(bits_of(n1), *gu_ck[1], g)
        want_mid = O.np_bf16_bits_to_f32(O.np_swiglu_bf16(O.np_f32_to_bf16_bits(gate.astype(np.float32)),
                                                          O.np_f32_to_bf16_bits(up.astype(np.float32)))).astype(np.float64)
        got_mid = O.np_bf16_bits_to_f32(bits_of(mid)).astype(np.float64)
        assert np.abs(got_mid - want_mid).max() <= 2e-2 * np.abs(want_mid).max()
        assert_close_ref(bits_of(dn), oracle_forward(bits_of(mid), *down_ck[:3], g), "down")
        assert_close_ref(bits_of(y_qkv), oracle_forward(bits_of(n2), *qkv_ck[:3], g), "qkv after input_layernorm")
        attn.copy_(t_bf16_from_bits(O.make_activations(M, H, seed=190 + M + rep)).to(DEV))   # next run: new inputs


def test_prologue_errors(product):
    lin, _ = make_linear(product, 4096, 256, 128, seed=5)
    x = torch.zeros((1, 4096), dtype=torch.bfloat16, device=DEV)
    y = torch.zeros((1, 256), dtype=torch.bfloat16, device=DEV)
    w = torch.ones(4096, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(product.W4A16Error, match="no earlier prologue"):
        product.DecodeChain([(lin, x, y)], prologues=[{"norm_weight": w, "add_running_residual": True}])
    with pytest.raises(product.W4A16Error, match="norm_weight"):
        product.DecodeChain([(lin, x, y)], prologues=[{"norm_weight": w[:128].contiguous()}])
    with pytest.raises(product.W4A16Error, match="one entry"):
        product.DecodeChain([(lin, x, y)], prologues=[])
    other, _ = make_linear(product, 4096, 256, 64, seed=6)           # mixed group sizes: the persistent kernel declines
    with pytest.raises(product.W4A16Error, match="persistent kernel"):
        product.DecodeChain([(lin, x, y), (other, x, torch.zeros((1, 256), dtype=torch.bfloat16, device=DEV))],
                            prologues=[{"norm_weight": w}, None])


def test_chain_of_gptq_formats(product):
    """The resident layout is format-agnostic: GPTQ with stored zeros and symmetric GPTQ (q - 8) linears in one chain
    with an AWQ one, each against its oracle."""
    from helpers import gptq_unpacked
    g, M = 128, 2
    qw, qz, sc, _ = O.make_gptq_checkpoint(4096, 512, g, seed=21)
    zp = product.GptqConfig(4, g, False).create_linear(4096, 512, False, DEV)
    zp.load_weights({"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc)})
    zp_ck = gptq_unpacked(qw, qz, sc)
    qw2, _, sc2, _ = O.make_gptq_checkpoint(4096, 768, g, seed=22)
    sym = product.MarlinConfig.gptq_int4(g).create_linear(4096, 768, False, DEV)
    sym.load_weights({"qweight": t_i32(qw2), "scales": torch.from_numpy(sc2)})
    sym_ck = gptq_unpacked(qw2, None, sc2)
    awq, awq_ck = make_linear(product, 4096, 256, g, seed=23)
    xs = [t_bf16_from_bits(O.make_activations(M, 4096, seed=30 + i)).to(DEV) for i in range(3)]
    ys = [torch.zeros((M, n), dtype=torch.bfloat16, device=DEV) for n in (512, 768, 256)]
    chain = product.DecodeChain([(zp, xs[0], ys[0]), (sym, xs[1], ys[1]), (awq, xs[2], ys[2])])
    assert chain.kernel().startswith("chain_imma")
    chain.run()
    torch.cuda.synchronize()
    assert_close_ref(bits_of(ys[0]), oracle_forward(bits_of(xs[0]), *zp_ck, g), "gptq zero-point")
    assert_close_ref(bits_of(ys[1]), oracle_forward(bits_of(xs[1]), sym_ck[0], None, sym_ck[2], g, zero_mode=1), "gptq symmetric")
    assert_close_ref(bits_of(ys[2]), oracle_forward(bits_of(xs[2]), *awq_ck[:3], g), "awq")


@pytest.mark.parametrize("shared", [False, True])
@pytest.mark.parametrize("M", [1, 3])
def test_act_order_linears_in_a_chain(product, M, shared):
    """GPTQ act-order (g_idx): the chain does the gather of x[:, perm] itself (BASELINE config 4 shapes, scaled down).
    Linear 0 reads the chain's input: the grid gathers it once into an ingest mailbox.  Linear 1 reads linear 0's output:
    alone (shared=False) its producer scatters into the mailbox in linear 1's order; together with a second act-order
    reader of the same output (shared=True) both gather from the shared mailbox into ingest mailboxes of their own."""
    from helpers import gptq_unpacked
    g = 128
    shapes = [(2560, 4096), (4096, 2560)] + ([(4096, 1024)] if shared else [])
    rng = np.random.default_rng(11)
    lins, cks = [], []
    for i, (K, N) in enumerate(shapes):
        qw, qz, sc, _ = O.make_gptq_checkpoint(K, N, g, seed=90 + i)
        sc = (sc * np.float16(0.3)).astype(np.float16)
        g_idx = rng.permutation(np.repeat(np.arange(K // g, dtype=np.int32), g)).astype(np.int32)
        lin = product.GptqConfig(4, g, True).create_linear(K, N, False, DEV)
        lin.load_weights({"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc), "g_idx": torch.from_numpy(g_idx)})
        lins.append(lin)
        cks.append(gptq_unpacked(qw, qz, sc) + (g_idx,))
    x0 = torch.zeros((M, 2560), dtype=torch.bfloat16, device=DEV)
    y0 = torch.zeros((M, 4096), dtype=torch.bfloat16, device=DEV)
    y1 = torch.zeros((M, 2560), dtype=torch.bfloat16, device=DEV)
    y2 = torch.zeros((M, 1024), dtype=torch.bfloat16, device=DEV)
    steps = [(lins[0], x0, y0), (lins[1], y0, y1)] + ([(lins[2], y0, y2)] if shared else [])
    chain = product.DecodeChain(steps)
    assert chain.kernel() == "chain_imma (1 mailboxes, 0 barriers)", chain.kernel()
    for rep in range(4):                       # both parities of every mailbox, twice
        x0.copy_(t_bf16_from_bits(O.make_activations(M, 2560, seed=120 + rep)).to(DEV))
        chain.run()
        torch.cuda.synchronize()
        outs = [(x0, y0), (y0, y1)] + ([(y0, y2)] if shared else [])
        for i, (xin, yout) in enumerate(outs):
            q, z, sb, gi = cks[i]
            assert_close_ref(bits_of(yout), oracle_forward(bits_of(xin), q, z, sb, g, g_idx=gi), f"act-order linear {i} rep {rep}")
