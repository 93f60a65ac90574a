"""CPU oracle for the W4A16 quantized-linear path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product (vllm-rust_b200/) must
never import it: a product path that routes through the oracle voids parity.

Two layers:
  * `C`  -- ctypes bindings to oracle/liboracle.so (w4a16_oracle.c), the scalar,
            loop-for-loop restatement of the reference (file:line cited there).
  * numpy mirrors (`np_*`) of the same arithmetic, vectorised, used for the
    larger parity cases.  tests/test_oracle_kat.py checks numpy == C bit-for-bit
    on small cases so the two cannot drift.

Parity status: pinned (reference KATs K-1..K-9 + reference CUDA kernels on the
GPU box); F1's GEMM order and g_idx act-order are unpinned, see w4a16_oracle.c.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

AWQ_UNDO_PACK = np.array([0, 4, 1, 5, 2, 6, 3, 7], dtype=np.uint32)  # awq.rs:38
AWQ_UNDO_PACK_SHIFTS = AWQ_UNDO_PACK * 4  # awq_marlin.rs:372


def build(force: bool = False) -> str:
    """Compile liboracle.so with gcc (seconds).  Building the checker is not using it."""
    src = os.path.join(_HERE, "w4a16_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def _ptr(a: np.ndarray | None):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "oracle wants contiguous arrays"
    return a.ctypes.data_as(ctypes.c_void_p)


class _C:
    """Lazy ctypes handle; every function takes/returns numpy arrays."""

    def __init__(self):
        self._lib = None

    @property
    def lib(self):
        if self._lib is None:
            build()
            lib = ctypes.CDLL(_LIB_PATH)
            sz, vp, u32, i32 = ctypes.c_size_t, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int
            lib.oracle_f32_to_bf16.restype = ctypes.c_uint16
            lib.oracle_f32_to_bf16.argtypes = [ctypes.c_float]
            lib.oracle_f32_to_f16.restype = ctypes.c_uint16
            lib.oracle_f32_to_f16.argtypes = [ctypes.c_float]
            lib.oracle_f16_to_f32.restype = ctypes.c_float
            lib.oracle_f16_to_f32.argtypes = [ctypes.c_uint16]
            lib.oracle_f32_to_bf16_array.argtypes = [vp, vp, sz]
            lib.oracle_f16_to_bf16_array.argtypes = [vp, vp, sz]
            lib.oracle_awq_unpack.argtypes = [vp, vp, sz, sz]
            lib.oracle_gptq_unpack_qweight.argtypes = [vp, vp, sz, sz]
            lib.oracle_seq_unpack.argtypes = [vp, vp, sz, sz]
            lib.oracle_awq_dequantize_cpu.argtypes = [vp, vp, vp, sz, sz, sz, vp, vp]
            lib.oracle_awq_linear_forward.argtypes = [vp, sz, vp, vp, vp, vp, sz, sz, sz, vp, vp, vp, i32]
            lib.oracle_awq_to_gptq_u32.restype = u32
            lib.oracle_awq_to_gptq_u32.argtypes = [u32]
            lib.oracle_repack_awq_nibbles.argtypes = [vp, vp, sz]
            lib.oracle_awq_to_gptq_qweight.restype = i32
            lib.oracle_awq_to_gptq_qweight.argtypes = [vp, vp, sz, sz]
            lib.oracle_transpose_u32.argtypes = [vp, vp, sz, sz]
            lib.oracle_awq_to_marlin_tile_repack.restype = i32
            lib.oracle_awq_to_marlin_tile_repack.argtypes = [vp, vp, sz, sz]
            lib.oracle_marlin_tile_decode_position.argtypes = [sz, sz, ctypes.POINTER(sz), ctypes.POINTER(u32)]
            lib.oracle_dequant_bf16.argtypes = [vp, vp, vp, vp, sz, sz, sz, i32, vp]
            lib.oracle_gemm_f32.argtypes = [vp, sz, vp, vp, vp, vp, vp, sz, sz, sz, i32, vp, i32]
            lib.oracle_lcg_words.argtypes = [u32, vp, sz]
            lib.oracle_lcg_gemm_case.argtypes = [u32, sz, sz, sz, sz, vp, vp, vp, vp]
            lib.oracle_max_threads.restype = i32
            self._lib = lib
        return self._lib

    # ---- unpack ----------------------------------------------------------
    def awq_unpack(self, packed: np.ndarray, cols: int) -> np.ndarray:
        rows = packed.shape[0]
        out = np.empty((rows, cols), np.uint8)
        self.lib.oracle_awq_unpack(_ptr(packed), _ptr(out), rows, cols)
        return out

    def gptq_unpack_qweight(self, qweight: np.ndarray, K: int, N: int) -> np.ndarray:
        out = np.empty((K, N), np.uint8)
        self.lib.oracle_gptq_unpack_qweight(_ptr(qweight), _ptr(out), K, N)
        return out

    def seq_unpack(self, packed: np.ndarray, cols: int) -> np.ndarray:
        rows = packed.shape[0]
        out = np.empty((rows, cols), np.uint8)
        self.lib.oracle_seq_unpack(_ptr(packed), _ptr(out), rows, cols)
        return out

    # ---- D1 / F1 ----------------------------------------------------------
    def awq_dequantize_cpu(self, qweight, scales_f32, qzeros, group_size, K, N) -> np.ndarray:
        """awq.rs:56-166 -> f16 [K,N] (returned as np.float16)."""
        scratch = np.empty((K, N), np.float32)
        out = np.empty((K, N), np.uint16)
        self.lib.oracle_awq_dequantize_cpu(_ptr(qweight), _ptr(scales_f32), _ptr(qzeros), group_size, K, N,
                                           _ptr(scratch), _ptr(out))
        return out.view(np.float16)

    def awq_linear_forward(self, x_bf16_bits, qweight, scales_f32, qzeros, bias_bf16_bits, group_size, K, N,
                           threads=1, scratch=None) -> np.ndarray:
        """awq.rs:522-567 -> bf16 bit patterns [M,N] (np.uint16)."""
        M = x_bf16_bits.shape[0]
        if scratch is None:
            scratch = (np.empty((K, N), np.float32), np.empty((M, K), np.float32))
        y = np.empty((M, N), np.uint16)
        self.lib.oracle_awq_linear_forward(_ptr(x_bf16_bits), M, _ptr(qweight), _ptr(scales_f32), _ptr(qzeros),
                                           _ptr(bias_bf16_bits), group_size, K, N, _ptr(scratch[0]),
                                           _ptr(scratch[1]), _ptr(y), threads)
        return y

    # ---- repacks -----------------------------------------------------------
    def awq_to_gptq_u32(self, w: int) -> int:
        return int(self.lib.oracle_awq_to_gptq_u32(ctypes.c_uint32(w)))

    def repack_awq_nibbles(self, words: np.ndarray) -> np.ndarray:
        out = np.empty_like(words)
        self.lib.oracle_repack_awq_nibbles(_ptr(words), _ptr(out), words.size)
        return out

    def awq_to_gptq_qweight(self, awq: np.ndarray, K: int, N: int) -> np.ndarray:
        out = np.empty((K // 8, N), np.uint32)
        rc = self.lib.oracle_awq_to_gptq_qweight(_ptr(awq), _ptr(out), K, N)
        if rc != 0:
            raise ValueError("awq_to_gptq_qweight: K and N must be divisible by 8")
        return out

    def transpose_u32(self, a: np.ndarray) -> np.ndarray:
        out = np.empty((a.shape[1], a.shape[0]), np.uint32)
        self.lib.oracle_transpose_u32(_ptr(a), _ptr(out), a.shape[0], a.shape[1])
        return out

    def awq_to_marlin_tile_repack(self, awq: np.ndarray, K: int, N: int) -> np.ndarray:
        if K % 16 or N % 64:
            raise ValueError("awq_to_marlin_tile_repack: size_k % 16 and size_n % 64 required")
        out = np.empty((K // 16, N * 2), np.uint32)
        self.lib.oracle_awq_to_marlin_tile_repack(_ptr(awq), _ptr(out), K, N)
        return out

    def marlin_tile_decode_position(self, k_in_tile: int, n_in_tile: int):
        a, b = ctypes.c_size_t(), ctypes.c_uint32()
        self.lib.oracle_marlin_tile_decode_position(k_in_tile, n_in_tile, ctypes.byref(a), ctypes.byref(b))
        return a.value, b.value

    # ---- D2 / GEMM ---------------------------------------------------------
    def dequant_bf16(self, q, z, scales_bf16_bits, g_idx, group_size, K, N, zero_mode=0) -> np.ndarray:
        out = np.empty((K, N), np.uint16)
        self.lib.oracle_dequant_bf16(_ptr(q), _ptr(z), _ptr(scales_bf16_bits), _ptr(g_idx), group_size, K, N,
                                     zero_mode, _ptr(out))
        return out

    def gemm_f32(self, x_f32, q, z, scales_f32, g_idx, bias_f32, group_size, K, N, zero_mode=0, threads=1):
        M = x_f32.shape[0]
        y = np.empty((M, N), np.float32)
        self.lib.oracle_gemm_f32(_ptr(x_f32), M, _ptr(q), _ptr(z), _ptr(scales_f32), _ptr(g_idx), _ptr(bias_f32),
                                 group_size, K, N, zero_mode, _ptr(y), threads)
        return y

    # ---- LCG ---------------------------------------------------------------
    def lcg_words(self, seed: int, n: int) -> np.ndarray:
        out = np.empty(n, np.uint32)
        self.lib.oracle_lcg_words(ctypes.c_uint32(seed), _ptr(out), n)
        return out

    def lcg_gemm_case(self, seed, M, K, N, group_size):
        """K-6 generator (marlin_cuda.rs:2054-2089): GPTQ-layout qweight [K/8,N],
        sequential qzeros [G,N/8], bf16-rounded f32 scales [G,N] and inputs [M,K]."""
        G = K // group_size
        qw = np.empty((K // 8, N), np.uint32)
        qz = np.empty((G, N // 8), np.uint32)
        sc = np.empty((G, N), np.float32)
        x = np.empty((M, K), np.float32)
        self.lib.oracle_lcg_gemm_case(ctypes.c_uint32(seed), M, K, N, group_size, _ptr(qw), _ptr(qz), _ptr(sc), _ptr(x))
        return qw, qz, sc, x

    def max_threads(self) -> int:
        return int(self.lib.oracle_max_threads())


C = _C()


# ---------------------------------------------------------------------------
# numpy mirrors (vectorised; same arithmetic, checked == C in the tests)
# ---------------------------------------------------------------------------

def np_f32_to_bf16_bits(a: np.ndarray) -> np.ndarray:
    """RNE f32 -> bf16 bit pattern (half::bf16::from_f32 / __float2bfloat16)."""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    nan = (u & 0x7FFFFFFF) > 0x7F800000
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    r[nan] = ((u[nan] >> 16) | 0x40).astype(np.uint16)
    return r


def np_bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b, np.uint16).astype(np.uint32) << 16).view(np.float32)


def np_bf16_round(a: np.ndarray) -> np.ndarray:
    return np_bf16_bits_to_f32(np_f32_to_bf16_bits(a))


def np_awq_unpack(packed: np.ndarray) -> np.ndarray:
    """[rows, cols/8] u32 -> [rows, cols] u8 with the AWQ interleave (awq.rs:38, 142-161)."""
    out = (packed[:, :, None] >> AWQ_UNDO_PACK_SHIFTS[None, None, :]) & 0xF
    return out.reshape(packed.shape[0], -1).astype(np.uint8)


def np_awq_pack(nib: np.ndarray) -> np.ndarray:
    """Inverse of np_awq_unpack: [rows, cols] u8 -> [rows, cols/8] u32."""
    r, c = nib.shape
    v = nib.reshape(r, c // 8, 8).astype(np.uint32)
    return (v << AWQ_UNDO_PACK_SHIFTS[None, None, :]).sum(axis=2, dtype=np.uint64).astype(np.uint32)


def np_seq_unpack(packed: np.ndarray) -> np.ndarray:
    """[rows, cols/8] u32 -> [rows, cols] u8, nibble i at bits 4i (GPTQ qzeros)."""
    sh = (np.arange(8, dtype=np.uint32) * 4)[None, None, :]
    return ((packed[:, :, None] >> sh) & 0xF).reshape(packed.shape[0], -1).astype(np.uint8)


def np_seq_pack(nib: np.ndarray) -> np.ndarray:
    r, c = nib.shape
    v = nib.reshape(r, c // 8, 8).astype(np.uint32)
    sh = (np.arange(8, dtype=np.uint32) * 4)[None, None, :]
    return (v << sh).sum(axis=2, dtype=np.uint64).astype(np.uint32)


def np_gptq_unpack_qweight(qweight: np.ndarray) -> np.ndarray:
    """[K/8, N] u32 -> [K, N] u8, nibble i of a word is k%8 (gptq_dequant.cu:76-80)."""
    kp, n = qweight.shape
    sh = (np.arange(8, dtype=np.uint32) * 4)[None, :, None]
    return ((qweight[:, None, :] >> sh) & 0xF).reshape(kp * 8, n).astype(np.uint8)


def np_gptq_pack_qweight(nib: np.ndarray) -> np.ndarray:
    k, n = nib.shape
    v = nib.reshape(k // 8, 8, n).astype(np.uint32)
    sh = (np.arange(8, dtype=np.uint32) * 4)[None, :, None]
    return (v << sh).sum(axis=1, dtype=np.uint64).astype(np.uint32)


def _group_index(K: int, group_size: int, g_idx: np.ndarray | None) -> np.ndarray:
    if g_idx is not None:
        return np.asarray(g_idx, np.int64)
    egs = K if group_size in (0, -1) else group_size
    return np.arange(K, dtype=np.int64) // egs


def np_dequant_bf16(q, z, scales_bf16_bits, group_size, g_idx=None, zero_mode=0) -> np.ndarray:
    """D2 (awq_marlin_dequant.cu:45-68 / gptq_dequant.cu:58-86): bf16 bit patterns [K,N]."""
    K, N = q.shape
    gi = _group_index(K, group_size, g_idx)
    s = np_bf16_bits_to_f32(scales_bf16_bits)[gi]
    zz = np.full((K, N), 8, np.int32) if zero_mode == 1 else z[gi].astype(np.int32)
    w = (q.astype(np.int32) - zz).astype(np.float32) * s
    return np_f32_to_bf16_bits(w)


def np_dequant_f32(q, z, scales_f32, group_size, g_idx=None, zero_mode=0) -> np.ndarray:
    """Un-rounded fp32 dequant: ((q - z) as f32) * scale (marlin_cuda.rs:2107)."""
    K, N = q.shape
    gi = _group_index(K, group_size, g_idx)
    zz = np.full((K, N), 8, np.int32) if zero_mode == 1 else z[gi].astype(np.int32)
    return (q.astype(np.int32) - zz).astype(np.float32) * np.ascontiguousarray(scales_f32, np.float32)[gi]


def np_gemm_ref(x_f32, q, z, scales_f32, group_size, g_idx=None, bias_f32=None, zero_mode=0) -> np.ndarray:
    """fp32 dequant-matmul for LARGE cases: same operands as oracle_gemm_f32 but
    accumulated in float64 (order-free), for the tolerance checks at full size."""
    w = np_dequant_f32(q, z, scales_f32, group_size, g_idx, zero_mode).astype(np.float64)
    y = np.ascontiguousarray(x_f32, np.float32).astype(np.float64) @ w
    if bias_f32 is not None:
        y = y + bias_f32.astype(np.float64)[None, :]
    return y


def np_swiglu_bf16(gate_bits: np.ndarray, up_bits: np.ndarray) -> np.ndarray:
    """silu(gate) * up on bf16 tensors, fp32 inside, one rounding to bf16: fused_swiglu_bf16 (swiglu.cu:17-19, 40-45)
    == silu_and_mul_separate_bf16 (activations.cu:34-36, 148-164), the kernel QuantizedSwiGluMlp::forward launches
    between gate_proj / up_proj and down_proj (qwen3_quantized.rs:532-551).  silu(x) = x / (1 + expf(-x)).
    The reference builds its kernels with --use_fast_math (build.rs:335-339: ex2.approx / rcp.approx), so the GPU
    result may differ from this exact-libm restatement in the last fp32 bits: after the bf16 rounding that is at most
    one bf16 ulp on a small fraction of elements (tests/test_fused_gpu.py pins the product bit-for-bit against the
    reference kernel itself and against this function within one ulp)."""
    g = np_bf16_bits_to_f32(gate_bits).astype(np.float32)
    u = np_bf16_bits_to_f32(up_bits).astype(np.float32)
    with np.errstate(over="ignore"):
        s = (g / (np.float32(1.0) + np.exp(-g, dtype=np.float32))).astype(np.float32)
    return np_f32_to_bf16_bits((s * u).astype(np.float32))


def reference_tolerance(expected: np.ndarray) -> float:
    """max(0.01 * max|expected|, 5e-2): marlin_cuda.rs:2168-2183, 2598-2611."""
    return max(0.01 * float(np.max(np.abs(expected))), 5e-2)


# ---------------------------------------------------------------------------
# synthetic checkpoints (BASELINE.md section 3 value distributions)
# ---------------------------------------------------------------------------

def make_awq_checkpoint(K: int, N: int, group_size: int = 128, seed: int = 0, bias: bool = False):
    """Uniform random nibbles and zeros, scales = f16(U(0.002, 0.02)); AWQ layout."""
    rng = np.random.default_rng(seed)
    G = K // group_size if group_size > 0 else 1
    qweight = rng.integers(0, 2**32, size=(K, N // 8), dtype=np.uint64).astype(np.uint32)
    qzeros = rng.integers(0, 2**32, size=(G, N // 8), dtype=np.uint64).astype(np.uint32)
    scales = rng.uniform(0.002, 0.02, size=(G, N)).astype(np.float16)
    b = rng.normal(0, 1, size=(N,)).astype(np.float32) if bias else None
    return qweight, qzeros, scales, b


def make_gptq_checkpoint(K: int, N: int, group_size: int = 128, seed: int = 0, act_order: bool = False):
    rng = np.random.default_rng(seed)
    G = K // group_size if group_size > 0 else 1
    qweight = rng.integers(0, 2**32, size=(K // 8, N), dtype=np.uint64).astype(np.uint32)
    qzeros = rng.integers(0, 2**32, size=(G, N // 8), dtype=np.uint64).astype(np.uint32)
    scales = rng.uniform(0.002, 0.02, size=(G, N)).astype(np.float16)
    g_idx = None
    if act_order:
        g_idx = rng.permutation(np.repeat(np.arange(G, dtype=np.int32), K // G)).astype(np.int32)
    return qweight, qzeros, scales, g_idx


def make_activations(M: int, K: int, seed: int = 1) -> np.ndarray:
    """x = bf16(N(0,1)) as bf16 bit patterns."""
    rng = np.random.default_rng(seed)
    return np_f32_to_bf16_bits(rng.normal(0, 1, size=(M, K)).astype(np.float32))


# ---- residual add + RMSNorm between the linears -------------------------------------------------------
def np_add_bf16(a_bits, b_bits):
    """add_bf16 (kernels/activations.cu:99-113): bf16_rn(f32(a) + f32(b)), elementwise."""
    return np_f32_to_bf16_bits(np_bf16_bits_to_f32(a_bits) + np_bf16_bits_to_f32(b_bits))


def np_rms_norm_bf16(x_bits, w_bits, eps):
    """rms_norm_bf16 (kernels/layernorm.cu:59-126): variance = sum(f32(x)^2) over the row (float64 here: the kernel's
    fp32 sum order is thread-strided + tree, so the last bit of inv_rms is not pinned), inv_rms = rsqrt(variance / H + eps)
    in fp32, out = bf16_rn((f32(x) * inv_rms) * f32(w)) with two fp32 multiplies in that order."""
    x = np_bf16_bits_to_f32(x_bits)
    w = np_bf16_bits_to_f32(w_bits)
    var = (x.astype(np.float64) ** 2).sum(axis=-1, keepdims=True)
    inv = (1.0 / np.sqrt(var / x.shape[-1] + np.float64(np.float32(eps)))).astype(np.float32)
    return np_f32_to_bf16_bits((x * inv).astype(np.float32) * w)
