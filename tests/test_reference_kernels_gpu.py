"""GPU parity against the REFERENCE'S OWN CUDA kernels (oracle/_ref/libref_kernels.so: the files under
/root/reference/crates/core/kernels compiled where they lie by `make -C oracle ref`, launched with the
geometry of marlin_cuda.rs / gptq_cuda.rs).  The reference's resident layouts are produced by the oracle's
restatement of its load-time repacks (awq_marlin.rs:580-667), so this checks, on the same inputs:
    product dequant   == reference awq_marlin_dequant_int4_bf16 / gptq_dequant_int4_bf16   bit for bit
    product forward   ~= reference awq_gemv_int4_kt_bf16 / gptq_gemm_int4_bf16_tiled        reference tolerance
The library is test infrastructure: it is never loaded by the product."""
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


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libref_kernels.so not built (needs /root/reference at build time)")
    lib = ctypes.CDLL(REF_SO)
    vp, i = ctypes.c_void_p, ctypes.c_int
    lib.ref_awq_marlin_dequant_int4_bf16.argtypes = [vp, vp, vp, vp, i, i, i, vp]
    lib.ref_awq_gemv_int4_kt_bf16.argtypes = [vp, vp, vp, vp, vp, vp, i, i, i, i, vp]
    lib.ref_gptq_dequant_int4_bf16.argtypes = [vp, vp, vp, vp, i, i, i, vp]
    lib.ref_gptq_gemm_int4_bf16_tiled.argtypes = [vp, vp, vp, vp, vp, vp, i, i, i, i, vp]
    return lib


def dev_i32(a):
    return t_i32(a).to(DEV)


def awq_resident(qweight, qzeros, scales, K, N):
    """AwqMarlinLinear::load_weights (awq_marlin.rs:621-656): [N,K/8] transposed sequential qweight,
    sequential qzeros, bf16 scales."""
    qw_t = O.C.transpose_u32(O.C.awq_to_gptq_qweight(qweight, K, N))
    qz = O.C.repack_awq_nibbles(qzeros)
    sb = O.np_f32_to_bf16_bits(np.asarray(scales, np.float32))
    return dev_i32(qw_t), dev_i32(qz), t_bf16_from_bits(sb).to(DEV)


@pytest.mark.parametrize("K,N,g", [(256, 128, 128), (512, 192, 64), (1024, 6144, 128), (4096, 4096, 128)])
def test_awq_dequant_matches_reference_kernel_bit_exact(product, ref, K, N, g):
    qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=K * 7 + N)
    lin = product.AwqConfig.int4(g).create_linear(K, N, False, DEV)
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    qw_t, qz, sb = awq_resident(qweight, qzeros, scales, K, N)
    out = torch.empty((K, N), dtype=torch.bfloat16, device=DEV)
    s = torch.cuda.current_stream().cuda_stream
    assert ref.ref_awq_marlin_dequant_int4_bf16(out.data_ptr(), qw_t.data_ptr(), sb.data_ptr(), qz.data_ptr(), K, N, g, s) == 0
    torch.cuda.synchronize()
    assert (bits_of(out) == bits_of(lin.dequantize())).all(), "product dequant != reference kernel"
    q, z, sbits = awq_unpacked(qweight, qzeros, scales)
    assert (bits_of(out) == O.np_dequant_bf16(q, z, sbits, g)).all(), "oracle dequant != reference kernel"


@pytest.mark.parametrize("M", [1, 4, 8])
@pytest.mark.parametrize("K,N", [(1024, 6144), (4096, 4096), (4096, 6144)])
def test_awq_gemv_matches_reference_kernel(product, ref, K, N, M):
    g = 128
    qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=K + N + M)
    lin = product.AwqConfig.int4(g).create_linear(K, N, False, DEV)
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    qw_t, qz, sb = awq_resident(qweight, qzeros, scales, K, N)
    xb = O.make_activations(M, K, seed=M)
    x = t_bf16_from_bits(xb).to(DEV)
    y_ref = torch.empty((M, N), dtype=torch.bfloat16, device=DEV)
    s = torch.cuda.current_stream().cuda_stream
    assert ref.ref_awq_gemv_int4_kt_bf16(y_ref.data_ptr(), x.data_ptr(), qw_t.data_ptr(), sb.data_ptr(), qz.data_ptr(),
                                         None, M, N, K, g, s) == 0
    y = lin.forward(x)
    torch.cuda.synchronize()
    q, z, sbits = awq_unpacked(qweight, qzeros, scales)
    want = oracle_forward(xb, q, z, sbits, g)
    assert_close_ref(bits_of(y_ref), want, "reference kernel vs oracle")   # pins the oracle to the reference
    assert_close_ref(bits_of(y), want, "product vs oracle")
    diff = np.abs(O.np_bf16_bits_to_f32(bits_of(y)) - O.np_bf16_bits_to_f32(bits_of(y_ref))).max()
    assert diff <= O.reference_tolerance(want)


@pytest.mark.parametrize("M", [1, 16, 64])
def test_gptq_matches_reference_kernels(product, ref, M):
    K, N, g = 2560, 1024, 128
    qweight, qzeros, scales, _ = O.make_gptq_checkpoint(K, N, g, seed=11 + M)
    lin = product.GptqConfig(4, g).create_linear(K, N, False, DEV)
    lin.load_weights({"qweight": t_i32(qweight), "qzeros": t_i32(qzeros), "scales": torch.from_numpy(scales)})
    sb = t_bf16_from_bits(O.np_f32_to_bf16_bits(scales.astype(np.float32))).to(DEV)
    qw, qz = dev_i32(qweight), dev_i32(qzeros)
    s = torch.cuda.current_stream().cuda_stream
    w_ref = torch.empty((K, N), dtype=torch.bfloat16, device=DEV)
    assert ref.ref_gptq_dequant_int4_bf16(w_ref.data_ptr(), qw.data_ptr(), sb.data_ptr(), qz.data_ptr(), K, N, g, s) == 0
    xb = O.make_activations(M, K, seed=M)
    x = t_bf16_from_bits(xb).to(DEV)
    y_ref = torch.empty((M, N), dtype=torch.bfloat16, device=DEV)
    assert ref.ref_gptq_gemm_int4_bf16_tiled(y_ref.data_ptr(), x.data_ptr(), qw.data_ptr(), sb.data_ptr(), qz.data_ptr(),
                                             None, M, N, K, g, s) == 0
    y = lin.forward(x)
    torch.cuda.synchronize()
    assert (bits_of(w_ref) == bits_of(lin.dequantize())).all(), "product GPTQ dequant != reference kernel"
    q, z, sbits = gptq_unpacked(qweight, qzeros, scales)
    want = oracle_forward(xb, q, z, sbits, g)
    assert_close_ref(bits_of(y_ref), want, "reference gptq_gemm vs oracle")
    assert_close_ref(bits_of(y), want, "product vs oracle")
