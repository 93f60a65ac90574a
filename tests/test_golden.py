"""Committed golden fixtures (tests/golden/w4a16_golden.npz, made by tests/golden/make_golden.py):
CPU: the oracle still reproduces them; GPU: the CUDA path, through the C ABI, matches them
(dequantised weights bit-exact, y within the reference tolerance)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
from helpers import assert_close_ref, bits_of, t_bf16_from_bits, t_i32

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "w4a16_golden.npz"))
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
MG = importlib.util.module_from_spec(spec)
spec.loader.exec_module(MG)


def test_oracle_reproduces_golden():
    fresh = MG.build()
    assert sorted(fresh) == sorted(GOLD.files)
    for k in GOLD.files:
        assert np.array_equal(np.asarray(fresh[k]), GOLD[k]), k


def test_golden_covers_every_case():
    for name in MG.CASES:
        assert name + "/y" in GOLD.files and name + "/dequant_checksum" in GOLD.files
    for name in MG.LCG:
        assert name + "/y" in GOLD.files


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(MG.CASES))
def test_cuda_path_matches_golden(product, name):
    kind, K, N, g, seed, M = MG.CASES[name]
    xb = O.make_activations(M, K, seed=seed + 100)
    if kind == "awq":
        qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=seed)
        lin = product.AwqConfig.int4(g).create_linear(K, N, False, "cuda:0")
        lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    else:
        qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, g, seed=seed, act_order=kind == "gptq_act")
        cfg = product.GptqConfig(4, g, desc_act=kind == "gptq_act", use_stored_zeros=kind != "gptq_sym")
        lin = cfg.create_linear(K, N, False, "cuda:0")
        w = {"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)}
        if g_idx is not None:
            w["g_idx"] = torch.from_numpy(g_idx)
        lin.load_weights(w)
    dq = bits_of(lin.dequantize())
    assert np.array_equal(MG.checksum(dq), GOLD[name + "/dequant_checksum"]), "dequantised weights differ from golden"
    if name + "/dequant_bf16_bits" in GOLD.files:
        assert np.array_equal(dq, GOLD[name + "/dequant_bf16_bits"])
    y = lin.forward(t_bf16_from_bits(xb).to("cuda:0"))
    torch.cuda.synchronize()
    assert_close_ref(bits_of(y), GOLD[name + "/y"].astype(np.float64), name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(MG.LCG))
def test_cuda_path_matches_lcg_golden(product, name):
    seed, M, K, N = MG.LCG[name]
    qw, qz, sc, x = O.C.lcg_gemm_case(seed, M, K, N, 128)
    lin = product.GptqConfig(4, 128).create_linear(K, N, False, "cuda:0")
    lin.load_weights({"qweight": t_i32(qw), "qzeros": t_i32(qz), "scales": torch.from_numpy(sc)})
    y = lin.forward(t_bf16_from_bits(O.np_f32_to_bf16_bits(x)).to("cuda:0"))
    torch.cuda.synchronize()
    assert_close_ref(bits_of(y), GOLD[name + "/y"].astype(np.float64), name)
