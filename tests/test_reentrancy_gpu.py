"""re-entrant for read-only weights.  Different handles on different streams at the same time -- including
the split-K prefill path, whose fp32 partials and tickets are per (device, stream) -- and the capture rules of the
per-handle scratch (never freed under a captured graph, never grown inside a capture)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import assert_close_ref, awq_unpacked, bits_of, gptq_unpacked, oracle_forward, t_bf16_from_bits, t_i32

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def awq_linear(product, K, N, g, seed):
    qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=seed)
    lin = product.AwqConfig.int4(g).create_linear(K, N, False, DEV)
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    return lin, awq_unpacked(qweight, qzeros, scales)


@pytest.mark.parametrize("M", [64, 1])
def test_two_streams_two_handles_concurrently(product, M):
    """M = 64 on 4096 -> 1024 / 2048 splits K over CTAs (few output tiles): each stream must use its own partials."""
    g = 128
    cases = [awq_linear(product, 4096, 1024, g, 1), awq_linear(product, 4096, 2048, g, 2)]
    xs = [t_bf16_from_bits(O.make_activations(M, 4096, seed=10 + i)).to(DEV) for i in range(2)]
    ys = [torch.empty((M, lin.out_features()), dtype=torch.bfloat16, device=DEV) for lin, _ in cases]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    for rep in range(20):                      # interleaved enqueues: the two streams overlap on the device
        for i, (lin, _) in enumerate(cases):
            with torch.cuda.stream(streams[i]):
                lin.forward_into(xs[i], ys[i])
    torch.cuda.synchronize()
    for i, (lin, (q, z, sb)) in enumerate(cases):
        assert_close_ref(bits_of(ys[i]), oracle_forward(bits_of(xs[i]), q, z, sb, g), f"stream {i} M={M}")
        with torch.cuda.stream(streams[i]):
            again = lin.forward(xs[i])
        torch.cuda.synchronize()
        assert (bits_of(again) == bits_of(ys[i])).all()      # deterministic, whatever ran beside it


def test_gather_scratch_is_not_grown_inside_a_capture(product):
    """Act-order forward needs a gather buffer sized for M: inside a capture that is an error naming w4a16_reserve;
    after reserve the capture works and replays; growing later does not invalidate the captured graph."""
    K, N, g = 1024, 256, 128
    qweight, qzeros, scales, _ = O.make_gptq_checkpoint(K, N, g, seed=3)
    g_idx = np.random.default_rng(0).permutation(np.repeat(np.arange(K // g, dtype=np.int32), g)).astype(np.int32)
    lin = product.GptqConfig(4, g, True).create_linear(K, N, False, DEV)
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales),
                      "g_idx": torch.from_numpy(g_idx)})
    q, z, sb = gptq_unpacked(qweight, qzeros, scales)
    x = t_bf16_from_bits(O.make_activations(8, K, seed=5)).to(DEV)
    y = torch.zeros((8, N), dtype=torch.bfloat16, device=DEV)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        graph = torch.cuda.CUDAGraph()
        with pytest.raises(product.W4A16Error, match="w4a16_reserve"):
            with torch.cuda.graph(graph, stream=st):
                lin.forward_into(x, y)
    torch.cuda.synchronize()
    lin.reserve(8)
    with torch.cuda.stream(st):
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            lin.forward_into(x, y)
        graph.replay()
    st.synchronize()
    want = oracle_forward(bits_of(x), q, z, sb, g, g_idx=g_idx)
    assert_close_ref(bits_of(y), want, "captured act-order forward")
    big = t_bf16_from_bits(O.make_activations(64, K, seed=6)).to(DEV)
    lin.forward(big)                               # grows the gather buffer; the old one must stay valid
    y.zero_()
    with torch.cuda.stream(st):
        graph.replay()
    st.synchronize()
    assert_close_ref(bits_of(y), want, "replay after the scratch grew")
