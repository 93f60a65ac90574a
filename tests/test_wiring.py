"""Wiring of the host-side mirror, ported from the reference's own integration tests
(/root/reference/tests/quantization_integration.rs:19-249): detection -> config -> weight loader ->
load_linear -> forward, for the W4A16 members of QuantizationMethod.  The detection / config / loader-selection
part runs on the CPU; the load_linear -> forward part needs the device (there is no CPU fallback)."""
import json

import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import assert_close_ref, awq_unpacked, bits_of, gptq_unpacked, oracle_forward, t_bf16_from_bits, t_i32


# ---- detection (quantization_integration.rs:19-38, detection.rs:70-262) ----------------------------------
def test_detect_from_nonexistent_directory_returns_none(product):
    assert product.detect_from_directory("/nonexistent/path").method is product.QuantizationMethod.NONE


def test_detect_from_directory_with_no_quant_config(product, tmp_path):
    (tmp_path / "config.json").write_text('{"architectures": ["LlamaForCausalLM"], "model_type": "llama"}')
    assert product.detect_from_directory(tmp_path).method is product.QuantizationMethod.NONE


@pytest.mark.parametrize("name,method", [("awq", "AWQ"), ("awq_triton", "AWQ"), ("awq_marlin", "AWQ_MARLIN"),
                                         ("gptq", "GPTQ"), ("gptq_marlin", "MARLIN"), ("fp8", "NONE")])
def test_detect_from_directory_config_json(product, tmp_path, name, method):
    cfg = {"model_type": "qwen3", "quantization_config": {"quant_method": name, "bits": 4, "group_size": 128,
                                                          "desc_act": True, "zero_point": True}}
    (tmp_path / "config.json").write_text(json.dumps(cfg))
    det = product.detect_from_directory(tmp_path)
    assert det.method is product.QuantizationMethod[method]   # methods off the W4A16 path are not this build's
    if method != "NONE":
        assert (det.bits, det.group_size, det.desc_act) == (4, 128, True)
        assert det.raw_config["zero_point"] is True


def test_detect_from_quantize_config_json(product, tmp_path):
    (tmp_path / "quantize_config.json").write_text('{"bits": 4, "group_size": 64, "desc_act": false}')
    det = product.detect_from_directory(tmp_path)
    assert (det.method, det.bits, det.group_size, det.desc_act) == (product.QuantizationMethod.GPTQ, 4, 64, False)
    (tmp_path / "quantize_config.json").write_text('{"bits": 4, "group_size": 128, "quant_method": "gptq_marlin"}')
    assert product.detect_from_directory(tmp_path).method is product.QuantizationMethod.MARLIN
    (tmp_path / "quantize_config.json").write_text('{"w_bit": 4, "q_group_size": 128, "zero_point": true}')
    det = product.detect_from_directory(tmp_path)
    assert (det.method, det.bits, det.group_size) == (product.QuantizationMethod.AWQ, 4, 128)


# ---- config creation (quantization_integration.rs:42-90) ---------------------------------------------------
def _detected(product, method, **kw):
    return product.DetectedQuantConfig(product.QuantizationMethod[method], kw.get("bits", 4), kw.get("group_size", 128),
                                       kw.get("desc_act"), kw.get("raw", {}))


def test_create_config_for_gptq(product):
    cfg = product.create_config(_detected(product, "GPTQ", desc_act=True))
    assert cfg.method() is product.QuantizationMethod.GPTQ
    assert (cfg.bits, cfg.group_size, cfg.desc_act) == (4, 128, True)


def test_create_config_for_awq_and_awq_marlin(product):
    cfg = product.create_config(_detected(product, "AWQ", raw={"modules_to_not_convert": ["lm_head", "mlp.gate"]}))
    assert cfg.method() is product.QuantizationMethod.AWQ
    assert cfg.is_layer_skipped("model.layers.0.mlp.gate") and not cfg.is_layer_skipped("model.layers.0.mlp.up_proj")
    cfgm = product.create_config(_detected(product, "AWQ_MARLIN"))
    assert cfgm.method() is product.QuantizationMethod.AWQ_MARLIN
    assert cfgm.is_layer_skipped("lm_head")                      # the default list (awq.rs:243-251)


def test_create_config_rejects_methods_off_the_path(product):
    with pytest.raises(product.W4A16Error, match="not on the W4A16 path"):
        product.create_config(product.DetectedQuantConfig())


def test_marlin_config_and_scalar_types(product):
    """marlin.rs:52-84, 88-266, 874-919."""
    g = product.MarlinConfig.gptq_int4(128)
    a = product.MarlinConfig.awq_int4(64)
    assert g.method() is product.QuantizationMethod.MARLIN and g.is_sym and not a.is_sym
    assert g.scalar_type.bits() == 4 and g.scalar_type.pack_factor() == 8 and not g.scalar_type.has_zero_point()
    assert a.scalar_type.has_zero_point()
    product.check_marlin_supports_shape(4096, 4096, 128)
    with pytest.raises(ValueError, match="GPTQ_MARLIN_MIN_THREAD_N"):
        product.check_marlin_supports_shape(100, 4096, 128)
    with pytest.raises(ValueError, match="GPTQ_MARLIN_MIN_THREAD_K"):
        product.check_marlin_supports_shape(4096, 100, 128)
    assert product.AwqConfig.int4(128).to_marlin_config().scalar_type is product.MarlinScalarType.UINT4
    assert product.AwqConfig(4, 100).to_marlin_config() is None   # group size outside the Marlin set


# ---- loader selection (quantization_integration.rs:94-146) ------------------------------------------------------
def test_create_weight_loader_for_gptq_and_awq(product):
    assert product.create_weight_loader_with_params({}, _detected(product, "GPTQ", desc_act=False), "cuda:0").method() \
        is product.QuantizationMethod.GPTQ
    assert product.create_weight_loader_with_params({}, _detected(product, "AWQ"), "cuda:0").method() \
        is product.QuantizationMethod.AWQ


def test_quantization_method_display_and_is_supported(product):
    """quantization_integration.rs:226-249."""
    QM = product.QuantizationMethod
    assert (str(QM.NONE), str(QM.GPTQ), str(QM.AWQ)) == ("none", "gptq", "awq")
    assert product.is_supported(70, QM.GPTQ) and not product.is_supported(60, QM.GPTQ)
    assert product.is_supported(80, QM.MARLIN) and not product.is_supported(70, QM.MARLIN)
    assert product.is_supported(80, QM.AWQ_MARLIN) and product.is_supported(0, QM.NONE)
    assert product.AwqConfig.int4(128).min_capability() == 100    # this build: sm_100a only


def test_mock_communicator_is_identity(product):
    """communicator.rs:100-236."""
    c = product.MockCommunicator(1, 4)
    t = torch.arange(6.0)
    assert (c.rank(), c.world_size(), c.nccl_comm()) == (1, 4, None)
    assert c.all_reduce(t) is t
    assert product.shard_range(4096, 3, 4) == (3072, 4096)
    with pytest.raises(ValueError):
        product.shard_range(100, 0, 3)


def test_loader_shape_and_missing_tensor_errors(product):
    """weight_loader.rs:626-671: a tensor that is missing or has the wrong shape fails with the prefix in the text."""
    ld = product.create_weight_loader_with_params({}, _detected(product, "AWQ"), "cuda:0")
    with pytest.raises(product.W4A16Error, match="qweight load failed for model.layers.0.self_attn.q_proj"):
        ld.load_linear("model.layers.0.self_attn.q_proj", 256, 128, False)
    bad = {"p.qweight": torch.zeros((256, 8), dtype=torch.int32)}
    ld = product.create_weight_loader_with_params(bad, _detected(product, "AWQ"), "cuda:0")
    with pytest.raises(product.W4A16Error, match=r"expected shape \[256, 16\]"):
        ld.load_linear("p", 256, 128, False)
    ld = product.create_weight_loader_with_params({}, _detected(product, "AWQ"), "cuda:0")
    with pytest.raises(product.W4A16Error, match="modules_to_not_convert"):
        ld.load_linear("lm_head", 256, 128, False)


# ---- detection -> config -> loader -> load_linear -> forward on the device ---------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("method", ["awq", "awq_marlin"])
def test_awq_wiring_end_to_end(product, method):
    K, N, g, M = 512, 256, 128, 3
    qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=11, bias=True)
    pre = "model.layers.0.self_attn.o_proj"
    tensors = {f"{pre}.qweight": t_i32(qweight), f"{pre}.qzeros": t_i32(qzeros),
               f"{pre}.scales": torch.from_numpy(scales), f"{pre}.bias": torch.from_numpy(b).to(torch.bfloat16)}
    det = product.detect_from_json({"quantization_config": {"quant_method": method, "bits": 4, "group_size": g,
                                                            "zero_point": True, "version": "gemm"}})
    loader = product.create_weight_loader_with_params(tensors, det, "cuda:0")
    lin = loader.load_linear(pre, K, N, True)
    assert (lin.in_features(), lin.out_features(), lin.has_bias(), lin.weight_dtype()) == (K, N, True, torch.int32)
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=5)
    y = lin.forward(t_bf16_from_bits(xb).to("cuda:0").view(1, M, K))        # rank-3 in, rank-3 out
    assert tuple(y.shape) == (1, M, N)
    assert_close_ref(bits_of(y).reshape(M, N), oracle_forward(xb, q, z, sb, g, bias_f32=O.np_bf16_round(b)), method)
    nobias = loader.load_linear(pre, K, N, False)                             # bias=false: the tensor is not read
    assert not nobias.has_bias()


@pytest.mark.gpu
@pytest.mark.parametrize("method,desc_act", [("gptq", False), ("gptq", True), ("gptq_marlin", False)])
def test_gptq_wiring_end_to_end(product, method, desc_act):
    K, N, g, M = 512, 128, 128, 2
    qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, g, seed=12, act_order=desc_act)
    pre = "model.layers.3.mlp.down_proj"
    tensors = {f"{pre}.qweight": t_i32(qweight), f"{pre}.qzeros": t_i32(qzeros), f"{pre}.scales": torch.from_numpy(scales)}
    if desc_act:
        tensors[f"{pre}.g_idx"] = torch.from_numpy(g_idx)
    det = product.detect_from_json({"quantization_config": {"quant_method": method, "bits": 4, "group_size": g,
                                                            "desc_act": desc_act, "sym": False}})
    loader = product.create_weight_loader_with_params(tensors, det, "cuda:0")
    assert loader.method() is product.QuantizationMethod.GPTQ
    lin = loader.load_linear(pre, K, N, False)
    q, z, sb = gptq_unpacked(qweight, qzeros, scales)
    xb = O.make_activations(M, K, seed=6)
    y = lin.forward(t_bf16_from_bits(xb).to("cuda:0"))
    assert_close_ref(bits_of(y), oracle_forward(xb, q, z, sb, g, g_idx=g_idx if desc_act else None), method)


@pytest.mark.gpu
def test_weight_groups_follow_the_constructor(product):
    """A B200W4A16Linear built directly (not through a config) reports its own group count."""
    K, N, g = 512, 128, 64
    qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=2)
    lin = product.B200W4A16Linear(K, N, False, g, 0, "cuda:0")
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    assert lin.weight_groups() == K // g
    q, z, sb = awq_unpacked(qweight, qzeros, scales)
    assert (lin.unpack()[1].cpu().numpy() == z).all() and (bits_of(lin.scales()) == sb).all()


@pytest.mark.gpu
def test_load_weights_rejects_a_checkpoint_of_another_group_size(product):
    K, N = 512, 128
    qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, 128, seed=2)       # scales [4, N]
    lin = product.AwqConfig.int4(64).create_linear(K, N, False, "cuda:0")        # expects [8, N]
    with pytest.raises(product.W4A16Error, match=r"scales shape .* expected \[8, 128\]"):
        lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, 128, seed=2, act_order=True)
    lin = product.GptqConfig(4, 128, desc_act=True).create_linear(K, N, False, "cuda:0")
    with pytest.raises(product.W4A16Error, match="g_idx must hold 512"):
        lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales),
                          "g_idx": torch.from_numpy(g_idx[:256])})
