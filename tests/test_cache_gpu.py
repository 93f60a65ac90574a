"""the on-disk cache of the resident layout (w4a16_save_tiled / w4a16_load_tiled) and the
safetensors -> resident-layout loader.  The reference repacks at every start (awq_marlin.rs:580-667).
A handle read back from the cache must be the same handle: hooks and forward outputs bit for bit; every header
field that does not match the receiving handle is an error that names it; corrupt files are refused."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import bits_of, t_bf16_from_bits, t_i32

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def awq_w(K, N, g, seed, bias=False):
    qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=seed, bias=bias)
    w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
    if bias:
        w["bias"] = torch.from_numpy(b).to(torch.bfloat16)
    return w


def same(a, b, K, Ms=(1, 4, 16, 200)):
    qa, za = a.unpack(); qb, zb = b.unpack()
    assert torch.equal(qa, qb) and torch.equal(za, zb)
    assert (bits_of(a.scales()) == bits_of(b.scales())).all()
    for M in Ms:
        x = t_bf16_from_bits(O.make_activations(M, K, seed=M)).to(DEV)
        assert (bits_of(a.forward(x)) == bits_of(b.forward(x))).all()


@pytest.mark.parametrize("fmt", ["awq", "awq_bias", "gptq_act_order", "column_shard", "swiglu"])
def test_cache_round_trip_is_bit_exact(product, tmp_path, fmt):
    K, N, g = 1024, 512, 128
    path = tmp_path / f"{fmt}.w4t"
    if fmt in ("awq", "awq_bias", "column_shard"):
        bias = fmt == "awq_bias"
        tp = dict(tp_mode=1, tp_rank=1, tp_size=2) if fmt == "column_shard" else {}
        make = lambda: product.AwqConfig.int4(g).create_linear(K, N, bias, DEV, **tp)   # noqa: E731
        a = make(); a.load_weights(awq_w(K, N, g, 3, bias))
    elif fmt == "gptq_act_order":
        qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, g, seed=4, act_order=True)
        make = lambda: product.GptqConfig(4, g, True).create_linear(K, N, False, DEV)   # noqa: E731
        a = make()
        a.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales),
                        "g_idx": torch.from_numpy(g_idx)})
    else:
        def make():
            lin = product.AwqConfig.int4(g).create_linear(K, N, False, DEV)
            lin.set_epilogue(product.EPILOGUE_SWIGLU)
            return lin
        a = make()
        a.load_part(0, awq_w(K, N // 2, g, 5)); a.load_part(N // 2, awq_w(K, N // 2, g, 6))
    a.save_tiled(path, fingerprint=0xABCDEF)
    assert not os.path.exists(str(path) + ".tmp")
    b = make()
    b.load_tiled(path, fingerprint=0xABCDEF)
    same(a, b, K)
    c = make()
    c.load_tiled(path)            # fingerprint 0: unchecked
    same(a, c, K, Ms=(2,))
    with pytest.raises(product.W4A16Error, match="fingerprint"):
        make().load_tiled(path, fingerprint=0x123)


def test_cache_refuses_mismatched_handles_and_corrupt_files(product, tmp_path):
    K, N, g = 512, 256, 128
    cfg = product.AwqConfig.int4(g)
    a = cfg.create_linear(K, N, False, DEV)
    with pytest.raises(product.W4A16Error, match="not loaded"):
        a.save_tiled(tmp_path / "x.w4t")
    a.load_weights(awq_w(K, N, g, 1))
    path = tmp_path / "a.w4t"
    a.save_tiled(path, 7)
    for make, field in [(lambda: cfg.create_linear(K, 2 * N, False, DEV), "N_full"),
                        (lambda: cfg.create_linear(2 * K, N, False, DEV), "K_full"),
                        (lambda: product.AwqConfig.int4(64).create_linear(K, N, False, DEV), "group_size_cfg"),
                        (lambda: cfg.create_linear(K, N, True, DEV), "has_bias"),
                        (lambda: product.GptqConfig.int4(g).create_linear(K, N, False, DEV), "format"),
                        (lambda: cfg.create_linear(K, N, False, DEV, tp_mode=2, tp_rank=0, tp_size=2), "K_local")]:
        with pytest.raises(product.W4A16Error, match=field):
            make().load_tiled(path, 7)
    sw = cfg.create_linear(K, N, False, DEV); sw.set_epilogue(product.EPILOGUE_SWIGLU)
    with pytest.raises(product.W4A16Error, match="epilogue"):
        sw.load_tiled(path, 7)
    raw = bytearray(path.read_bytes())
    # flipped payload bit -> checksum
    bad = bytearray(raw); bad[200] ^= 1
    (tmp_path / "bit.w4t").write_bytes(bad)
    with pytest.raises(product.W4A16Error, match="checksum"):
        cfg.create_linear(K, N, False, DEV).load_tiled(tmp_path / "bit.w4t", 7)
    (tmp_path / "short.w4t").write_bytes(raw[:-16])
    with pytest.raises(product.W4A16Error, match="payload size"):
        cfg.create_linear(K, N, False, DEV).load_tiled(tmp_path / "short.w4t", 7)
    (tmp_path / "long.w4t").write_bytes(raw + b"\0")
    with pytest.raises(product.W4A16Error, match="payload size"):
        cfg.create_linear(K, N, False, DEV).load_tiled(tmp_path / "long.w4t", 7)
    (tmp_path / "magic.w4t").write_bytes(b"NOTACACH" + raw[8:])
    with pytest.raises(product.W4A16Error, match="magic"):
        cfg.create_linear(K, N, False, DEV).load_tiled(tmp_path / "magic.w4t", 7)
    (tmp_path / "tiny.w4t").write_bytes(raw[:40])
    with pytest.raises(product.W4A16Error, match="shorter than the header"):
        cfg.create_linear(K, N, False, DEV).load_tiled(tmp_path / "tiny.w4t", 7)
    with pytest.raises(product.W4A16Error, match="cannot open"):
        cfg.create_linear(K, N, False, DEV).load_tiled(tmp_path / "missing.w4t", 7)
    # a failed load leaves the handle unloaded
    fresh = cfg.create_linear(K, N, False, DEV)
    with pytest.raises(product.W4A16Error):
        fresh.load_tiled(tmp_path / "bit.w4t", 7)
    with pytest.raises(product.W4A16Error, match="not loaded"):
        fresh.forward(torch.zeros((1, K), dtype=torch.bfloat16, device=DEV))


def tiny_qwen3_dir(product, tmp_path, skip=(), shards=1, layers=2):
    """A Qwen3-shaped AWQ checkpoint directory: config.json with quantization_config + safetensors."""
    H, I, heads, kv, hd, g = 512, 1024, 8, 4, 64, 128
    cfg = {"model_type": "qwen3", "hidden_size": H, "intermediate_size": I, "num_attention_heads": heads,
           "num_key_value_heads": kv, "head_dim": hd, "num_hidden_layers": layers,
           "quantization_config": {"quant_method": "awq", "bits": 4, "group_size": g, "zero_point": True,
                                   "version": "gemm", "modules_to_not_convert": ["lm_head", *skip]}}
    tmp_path.mkdir(parents=True, exist_ok=True)
    (tmp_path / "config.json").write_text(json.dumps(cfg))
    shapes = {"self_attn.q_proj": (H, heads * hd), "self_attn.k_proj": (H, kv * hd), "self_attn.v_proj": (H, kv * hd),
              "self_attn.o_proj": (heads * hd, H), "mlp.gate_proj": (H, I), "mlp.up_proj": (H, I), "mlp.down_proj": (I, H)}
    tensors, seed = {}, 0
    for li in range(layers):
        for mod, (k, n) in shapes.items():
            prefix = f"model.layers.{li}.{mod}"
            seed += 1
            if any(s in prefix for s in skip):
                tensors[f"{prefix}.weight"] = torch.randn(n, k, generator=torch.Generator().manual_seed(seed)).to(torch.bfloat16)
                continue
            for name, t in awq_w(k, n, g, seed).items():
                tensors[f"{prefix}.{name}"] = t
    tensors["lm_head.weight"] = torch.zeros(8, H, dtype=torch.bfloat16)
    if shards == 1:
        product.save_safetensors(tensors, tmp_path / "model.safetensors", {"format": "pt"})
    else:
        product.save_sharded(tensors, tmp_path, shards)
    return tensors, shapes, g


@pytest.mark.parametrize("shards", [1, 3])
def test_safetensors_directory_to_resident_layout_with_cache(product, tmp_path, shards):
    model = tmp_path / "model"
    tensors, shapes, g = tiny_qwen3_dir(product, model, skip=("layers.1.mlp.down_proj",), shards=shards)
    cache = tmp_path / "cache"
    layers, dense, loader = product.load_decoder_linears(model, DEV, cache_dir=cache)
    assert loader.cache_hits == 0
    assert sorted(layers[0]) == ["down_proj", "gate_up", "o_proj", "qkv"]
    assert sorted(layers[1]) == ["gate_up", "o_proj", "qkv"]                      # down_proj of layer 1 is dense
    assert list(dense) == ["model.layers.1.mlp.down_proj"]
    assert torch.equal(dense["model.layers.1.mlp.down_proj"], tensors["model.layers.1.mlp.down_proj.weight"])
    assert len(list(cache.iterdir())) == 7
    # second start: everything comes from the cache, and is the same
    layers2, _, loader2 = product.load_decoder_linears(model, DEV, cache_dir=cache)
    assert loader2.cache_hits == 7
    H = 512
    x = t_bf16_from_bits(O.make_activations(3, H, seed=1)).to(DEV)
    for li in layers:
        for name in layers[li]:
            a, b = layers[li][name], layers2[li][name]
            xin = x if a.in_features() == H else t_bf16_from_bits(O.make_activations(3, a.in_features(), seed=2)).to(DEV)
            assert (bits_of(a.forward(xin)) == bits_of(b.forward(xin))).all(), (li, name)
    # unfused == the reference's seven linears; fused qkv == their concatenation
    seven, _, _ = product.load_decoder_linears(model, DEV, layers=[0], fuse=False)
    assert sorted(seven[0]) == ["down_proj", "gate_proj", "k_proj", "o_proj", "q_proj", "up_proj", "v_proj"]
    q, k, v = (seven[0][n].forward(x) for n in ("q_proj", "k_proj", "v_proj"))
    got = layers[0]["qkv"].forward(x)
    want = torch.cat([q, k, v], 1)
    assert got.shape == want.shape
    assert (got.float() - want.float()).abs().max() <= 1e-2 * want.float().abs().max()
    assert layers[0]["gate_up"].out_features() == 1024
    # a changed checkpoint invalidates the cache (fingerprint), it is rebuilt, not trusted
    tensors2, _, _ = tiny_qwen3_dir(product, model, skip=("layers.1.mlp.down_proj",), shards=shards, layers=3)
    _, _, loader3 = product.load_decoder_linears(model, DEV, layers=[0], cache_dir=cache)
    assert loader3.cache_hits == 0


def test_skipped_module_is_not_on_the_w4a16_path(product, tmp_path):
    """modules_to_not_convert (awq.rs:243-251): load_linear refuses, the directory loader hands the dense weight back."""
    model = tmp_path / "m"
    tiny_qwen3_dir(product, model, skip=("self_attn.k_proj",), layers=1)
    layers, dense, loader = product.load_decoder_linears(model, DEV)
    assert "qkv" not in layers[0] and sorted(layers[0]) == ["down_proj", "gate_up", "o_proj", "q_proj", "v_proj"]
    assert list(dense) == ["model.layers.0.self_attn.k_proj"]
    with pytest.raises(product.W4A16Error, match="modules_to_not_convert"):
        loader.load_linear("model.layers.0.self_attn.k_proj", 512, 256, False)
    with pytest.raises(product.W4A16Error, match="modules_to_not_convert"):
        loader.load_fused(["model.layers.0.self_attn.q_proj", "model.layers.0.self_attn.k_proj"], 512, [512, 256], False)
