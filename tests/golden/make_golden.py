#!/usr/bin/env python
"""Regenerates tests/golden/w4a16_golden.npz from the CPU oracle (oracle/w4a16_oracle.c + oracle.py).

The reference is a Rust crate that cannot be built in this image (no cargo, crates.io dependencies not
vendored), so the fixtures are outputs of the oracle -- which tests/test_oracle_kat.py pins to the
reference's literal known-answer values (awq_marlin.rs:727-799, marlin_cuda.rs:1941-2112) and which
tests/test_reference_kernels_gpu.py pins to the reference's own CUDA kernels on the GPU box.
Inputs are seeded generators (numpy default_rng / the reference's LCG), so only seeds and outputs are stored.
    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle as O  # noqa: E402
from helpers import awq_unpacked, gptq_unpacked, oracle_forward  # noqa: E402

CASES = {
    # name: (kind, K, N, g, seed, M)
    "awq_256x128_g128": ("awq", 256, 128, 128, 0, 3),
    "awq_512x192_g64": ("awq", 512, 192, 64, 1, 1),
    "awq_cfg1_1024x6144_g128": ("awq", 1024, 6144, 128, 2, 1),   # BASELINE config 1
    "gptq_256x64_g128_actorder": ("gptq_act", 256, 64, 128, 3, 2),
    "gptq_256x64_g128_sym": ("gptq_sym", 256, 64, 128, 4, 2),
}
LCG = {"lcg_m1": (0x9E3779B9, 1, 256, 128), "lcg_m4": (0xCAFEBABE, 4, 256, 64), "lcg_m32": (0xDEADBEEF, 32, 256, 64)}


def checksum(bits):
    """Position-sensitive checksum of a uint16 array: (sum of bits * (index % 65521 + 1), xor) as uint64."""
    b = np.ascontiguousarray(bits).ravel().astype(np.uint64)
    w = (np.arange(b.size, dtype=np.uint64) % np.uint64(65521)) + np.uint64(1)
    return np.array([np.sum(b * w, dtype=np.uint64), np.bitwise_xor.reduce(b)], np.uint64)


def build():
    O.build()
    out = {}
    for name, (kind, K, N, g, seed, M) in CASES.items():
        xb = O.make_activations(M, K, seed=seed + 100)
        if kind == "awq":
            qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=seed)
            q, z, sb = awq_unpacked(qweight, qzeros, scales)
            dq = O.np_dequant_bf16(q, z, sb, g)
            if K * N <= 512 * 192:
                out[name + "/dequant_bf16_bits"] = dq
            out[name + "/dequant_checksum"] = checksum(dq)
            out[name + "/y"] = oracle_forward(xb, q, z, sb, g).astype(np.float32)
            if K * N <= 512 * 192:   # D1 grid of AwqLinear (awq.rs:56-166): f16 scales, f16 result
                d1 = O.C.awq_dequantize_cpu(qweight, scales.astype(np.float32), qzeros, g, K, N)
                out[name + "/d1_f16_bits"] = np.asarray(d1).view(np.uint16) if d1.dtype == np.float16 else np.asarray(d1)
        else:
            qweight, qzeros, scales, g_idx = O.make_gptq_checkpoint(K, N, g, seed=seed, act_order=kind == "gptq_act")
            q, z, sb = gptq_unpacked(qweight, qzeros, scales)
            zm = 1 if kind == "gptq_sym" else 0
            out[name + "/dequant_bf16_bits"] = O.np_dequant_bf16(q, z, sb, g, g_idx=g_idx, zero_mode=zm)
            out[name + "/dequant_checksum"] = checksum(out[name + "/dequant_bf16_bits"])
            out[name + "/y"] = oracle_forward(xb, q, z if zm == 0 else None, sb, g, g_idx=g_idx, zero_mode=zm).astype(np.float32)
    for name, (seed, M, K, N) in LCG.items():   # marlin_cuda.rs:2032-2612 generators, scalar fp32 oracle
        qw, qz, sc, x = O.C.lcg_gemm_case(seed, M, K, N, 128)
        q, z = O.C.gptq_unpack_qweight(qw, K, N), O.C.seq_unpack(qz, N)
        out[name + "/y"] = np.asarray(O.C.gemm_f32(x, q, z, sc, None, None, 128, K, N), np.float32)
        out[name + "/qweight_crc"] = np.array([int(np.bitwise_xor.reduce(qw.ravel()))], np.uint64)
    return out


if __name__ == "__main__":
    data = build()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "w4a16_golden.npz")
    np.savez_compressed(path, **data)
    print("wrote", path, os.path.getsize(path), "bytes,", len(data), "arrays")
