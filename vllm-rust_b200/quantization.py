"""Host-side mirror of vllm-rust's W4A16 quantization API over libw4a16_b200.so.

Reference interfaces mirrored (paths under crates/core/src/quantization/):
  trait QuantizationConfig          config.rs:110-135
  trait QuantizedLinear             config.rs:147-190
  AwqConfig / AwqVersion            awq.rs:170-393
  AwqMarlinConfig                   awq_marlin.rs:38-114
  MarlinConfig / MarlinScalarType   marlin.rs:52-266
  GptqConfig                        gptq.rs:28-246
  check_marlin_supports_shape       marlin.rs:308-341
  AwqWeightLoader / GptqWeightLoader / create_weight_loader_with_params
                                    weight_loader.rs:253-317, 596-714, 1596-1706
  detect_from_json                  detection.rs:240-262

Same names, argument meaning and error behaviour; the tensors are torch CUDA
tensors (device memory + streams only -- every computation is a kernel of the
library).  There is no CPU fallback: shapes outside the Marlin gate, which the
reference routes to its CPU-dequant AwqLinear (awq.rs:384-393), raise here.
"""
from __future__ import annotations

import ctypes
import enum
import hashlib
import os
from dataclasses import dataclass, field
from typing import Dict, Mapping, Optional, Sequence

import torch

from ._lib import W4A16Error, check, lib

GPTQ_MARLIN_MIN_THREAD_N = 64   # marlin.rs:41
GPTQ_MARLIN_MIN_THREAD_K = 128  # marlin.rs:43
MARLIN_SUPPORTED_GROUP_SIZES = (-1, 32, 64, 128)  # marlin.rs:48

FORMAT_AWQ, FORMAT_GPTQ_ZP, FORMAT_GPTQ_SYM = 0, 1, 2
DTYPE_CODE = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}
TP_NONE, TP_COLUMN, TP_ROW = 0, 1, 2


EPILOGUE_NONE, EPILOGUE_SWIGLU = 0, 1   # enum w4a16_epilogue (include/w4a16_b200.h)


class QuantizationMethod(enum.Enum):  # quantization/config.rs QuantizationMethod (the W4A16 members)
    NONE = "none"
    AWQ = "awq"
    AWQ_MARLIN = "awq_marlin"
    GPTQ = "gptq"
    MARLIN = "marlin"  # "gptq_marlin" routes here (detection.rs:258)

    def __str__(self) -> str:  # Display impl (config.rs): the lower-case method name
        return self.value


class MarlinScalarType(enum.Enum):  # marlin.rs:52-84
    UINT4B8 = "uint4b8"  # symmetric, W = (q - 8) * s
    UINT4 = "uint4"      # runtime zero point (AWQ)

    def bits(self) -> int:
        return 4

    def has_zero_point(self) -> bool:
        return self is MarlinScalarType.UINT4

    def pack_factor(self) -> int:
        return 8


def check_marlin_supports_shape(output_size: int, input_size: int, group_size: int) -> None:
    """marlin.rs:308-341.  Raises ValueError with the reference's wording."""
    if output_size % GPTQ_MARLIN_MIN_THREAD_N:
        raise ValueError(f"Marlin requires output_size ({output_size}) to be divisible by "
                         f"{GPTQ_MARLIN_MIN_THREAD_N} (GPTQ_MARLIN_MIN_THREAD_N)")
    if input_size % GPTQ_MARLIN_MIN_THREAD_K:
        raise ValueError(f"Marlin requires input_size ({input_size}) to be divisible by "
                         f"{GPTQ_MARLIN_MIN_THREAD_K} (GPTQ_MARLIN_MIN_THREAD_K)")
    if group_size > 0 and input_size % group_size:
        raise ValueError(f"Marlin requires input_size ({input_size}) to be divisible by group_size ({group_size})")


class QuantizedLinear:
    """config.rs:147-190."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def load_weights(self, weights: Mapping[str, torch.Tensor]) -> None:
        raise NotImplementedError

    def weight_dtype(self) -> torch.dtype:
        raise NotImplementedError

    def in_features(self) -> int:
        raise NotImplementedError

    def out_features(self) -> int:
        raise NotImplementedError

    def has_bias(self) -> bool:
        raise NotImplementedError


class QuantizationConfig:
    """config.rs:110-135."""

    def method(self) -> QuantizationMethod:
        raise NotImplementedError

    def supported_act_dtypes(self) -> Sequence[torch.dtype]:
        return (torch.bfloat16,)

    def min_capability(self) -> int:
        return 100  # sm_100a only

    def is_layer_skipped(self, layer_name: str) -> bool:
        return False

    def create_linear(self, in_features: int, out_features: int, bias: bool, device) -> QuantizedLinear:
        raise NotImplementedError


def _stream_ptr(x: Optional[torch.Tensor] = None) -> int:
    return torch.cuda.current_stream(x.device if x is not None else None).cuda_stream


def _device_index(device) -> int:
    d = torch.device(device)
    if d.type != "cuda":
        raise W4A16Error(f"B200W4A16Linear needs a CUDA device, got '{d}' (there is no CPU fallback)")
    return d.index if d.index is not None else torch.cuda.current_device()


class B200W4A16Linear(QuantizedLinear):
    """The drop-in for AwqMarlinLinear (awq_marlin.rs:431-667), MarlinLinear
    (marlin.rs:586-778) and GptqLinear (gptq.rs:254-471) on sm_100a."""

    def __init__(self, in_features: int, out_features: int, bias: bool, group_size: int, fmt: int, device,
                 tp_mode: int = TP_NONE, tp_rank: int = 0, tp_size: int = 1):
        self._h = ctypes.c_void_p()
        self._device = torch.device("cuda", _device_index(device))
        check(lib().w4a16_create_shard(ctypes.byref(self._h), in_features, out_features, group_size, fmt,
                                       int(bias), self._device.index, tp_mode, tp_rank, tp_size))
        self._full = (in_features, out_features)
        self._fmt = fmt
        self._group_size = group_size
        self._loaded = False

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                lib().w4a16_destroy(h)
            except Exception:
                pass
            self._h = None

    # -- trait surface ---------------------------------------------------------------
    def in_features(self) -> int:
        return int(lib().w4a16_in_features(self._h))

    def out_features(self) -> int:
        return int(lib().w4a16_out_features(self._h))

    def has_bias(self) -> bool:
        return bool(lib().w4a16_has_bias(self._h))

    def weight_dtype(self) -> torch.dtype:
        return torch.int32  # DType::U32 in the reference (awq_marlin.rs weight_dtype); torch has no stable u32

    def weight_bytes(self) -> int:
        return int(lib().w4a16_weight_bytes(self._h))

    def kernel_for(self, m: int) -> str:
        return lib().w4a16_kernel_for(self._h, m).decode()

    def load_weights(self, weights: Mapping[str, torch.Tensor]) -> None:
        """Keys "qweight", "qzeros", "scales", "bias", "g_idx" in CHECKPOINT layout
        (awq_marlin.rs:592-664, gptq.rs:435-454); the repack runs inside."""
        self._load(weights, None, self._full[1])

    def load_part(self, n_begin: int, weights: Mapping[str, torch.Tensor]) -> None:
        """One sibling projection of a fused handle (w4a16_load_*_part): `weights` is that projection's own
        checkpoint (the keys load_weights takes), filling output columns [n_begin, n_begin + its width).  The
        reference keeps q/k/v and gate/up as separate linears (qwen3_quantized.rs:84-96, 532-551)."""
        scales = weights.get("scales")
        if scales is None:
            raise W4A16Error("Missing 'scales' in weights")
        self._load(weights, int(n_begin), int(scales.shape[-1]))

    def set_epilogue(self, epilogue: int) -> None:
        """EPILOGUE_SWIGLU: the handle holds gate_proj | up_proj and forward writes silu(gate) * up
        (fused_swiglu_bf16, swiglu.cu:40-45); before any load."""
        check(lib().w4a16_set_epilogue(self._h, int(epilogue)))

    def _load(self, weights: Mapping[str, torch.Tensor], n_begin: Optional[int], nf: int) -> None:
        def need(key):
            t = weights.get(key)
            if t is None:
                raise W4A16Error(f"Missing '{key}' in weights")  # awq_marlin.rs:593-598 wording
            return t

        def dev(t, dtype=None):
            if t is None:
                return None
            if dtype is not None and t.dtype != dtype:
                if t.dtype.itemsize != 4 or t.dtype.is_floating_point:
                    raise W4A16Error(f"packed tensor must be a 32-bit integer type, got {t.dtype}")
                t = t.view(dtype)  # checkpoints hold the packed words as I32 / U32: same bits
            return t.to(self._device).contiguous()

        kf = self._full[0]
        qweight = dev(need("qweight"), torch.int32)
        scales = dev(need("scales"))
        if scales.dtype not in DTYPE_CODE:
            raise W4A16Error(f"scales dtype {scales.dtype} not supported")
        qzeros = dev(weights.get("qzeros"), torch.int32)
        bias = dev(weights.get("bias"))
        if bias is not None and bias.dtype not in DTYPE_CODE:
            raise W4A16Error(f"bias dtype {bias.dtype} not supported")
        if bias is not None and bias.numel() != nf:
            raise W4A16Error(f"bias holds {bias.numel()} entries, expected {nf}")
        g = 1 if self._group_size <= 0 else kf // self._group_size   # marlin.rs:308-341: K % group_size == 0
        exp_qw = (kf, nf // 8) if self._fmt == FORMAT_AWQ else (kf // 8, nf)
        if tuple(qweight.shape) != exp_qw:
            raise W4A16Error(f"qweight shape {tuple(qweight.shape)} != expected {exp_qw}")
        if tuple(scales.shape) != (g, nf):
            raise W4A16Error(f"scales shape {tuple(scales.shape)} != expected [{g}, {nf}]")
        if qzeros is not None and tuple(qzeros.shape) != (g, nf // 8):
            raise W4A16Error(f"qzeros shape {tuple(qzeros.shape)} != expected [{g}, {nf // 8}]")
        ptr = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
        bias_code = DTYPE_CODE[bias.dtype] if bias is not None else 1
        part = () if n_begin is None else (n_begin, nf)
        with torch.cuda.device(self._device):
            s = _stream_ptr(qweight)
            if self._fmt == FORMAT_AWQ:
                if qzeros is None:
                    raise W4A16Error("Missing 'qzeros' in weights")
                fn = lib().w4a16_load_awq if n_begin is None else lib().w4a16_load_awq_part
                check(fn(self._h, *part, ptr(qweight), ptr(qzeros), ptr(scales), DTYPE_CODE[scales.dtype],
                         ptr(bias), bias_code, s))
            else:
                g_idx = weights.get("g_idx")
                if g_idx is not None:
                    if g_idx.dtype.is_floating_point or g_idx.numel() != kf:
                        raise W4A16Error(f"g_idx must hold {kf} integer entries, got {g_idx.numel()} of {g_idx.dtype}")
                    g_idx = g_idx.to(self._device, torch.int32).contiguous()
                fn = lib().w4a16_load_gptq if n_begin is None else lib().w4a16_load_gptq_part
                check(fn(self._h, *part, ptr(qweight), ptr(qzeros), ptr(scales), DTYPE_CODE[scales.dtype], ptr(g_idx),
                         ptr(bias), bias_code, s))
            # the checkpoint tensors are locals: keep them alive until the repack has run
            torch.cuda.current_stream().synchronize()
        self._loaded = True

    # -- cache of the resident layout (w4a16_save_tiled / w4a16_load_tiled) -------------------------
    def save_tiled(self, path, fingerprint: int = 0) -> None:
        check(lib().w4a16_save_tiled(self._h, os.fsencode(path), fingerprint & (2 ** 64 - 1)))

    def load_tiled(self, path, fingerprint: int = 0) -> None:
        """Replaces load_weights: reads the repacked bytes back with plain copies; the handle must have been created
        with the geometry, format, bias flag, shard and epilogue the file was written from."""
        with torch.cuda.device(self._device):
            check(lib().w4a16_load_tiled(self._h, os.fsencode(path), fingerprint & (2 ** 64 - 1), _stream_ptr()))
        self._loaded = True

    def weight_rows(self) -> int:
        return int(lib().w4a16_weight_rows(self._h))

    def _check_x(self, x: torch.Tensor):
        if not x.is_cuda:
            raise W4A16Error("forward: input must be a CUDA tensor (there is no CPU fallback)")
        if x.dtype != torch.bfloat16:
            # marlin_cuda.rs:138-146, 492-495: the GPU path takes BF16 only
            raise W4A16Error(f"forward: input dtype must be BF16, got {x.dtype}")
        if x.shape[-1] != self.in_features():
            raise W4A16Error(f"forward: input last dim {x.shape[-1]} != in_features {self.in_features()}")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """config.rs:149.  Rank-2 or rank-3 input, leading dims flattened and restored
        (awq_marlin.rs:545-571); x made contiguous first (marlin_cuda.rs:858-869)."""
        self._check_x(x)
        lead = x.shape[:-1]
        x2 = x.contiguous().view(-1, x.shape[-1])
        y = torch.empty((x2.shape[0], self.out_features()), dtype=torch.bfloat16, device=x.device)
        self.forward_into(x2, y)
        return y.view(*lead, self.out_features())

    def forward_into(self, x2: torch.Tensor, y: torch.Tensor) -> None:
        """Caller-owned output, no allocation, graph-capturable: the OutputPool contract
        of marlin_gemm_pooled (marlin_cuda.rs:1655-1753) / forward_pooled (config.rs:167-174)."""
        m = x2.shape[0]
        if m < 1:
            raise W4A16Error("forward: M must be >= 1")
        check(lib().w4a16_forward_into(self._h, ctypes.c_void_p(x2.data_ptr()), ctypes.c_void_p(y.data_ptr()), m,
                                       _stream_ptr(x2)))

    def forward_host(self, x_host: torch.Tensor, y_host: torch.Tensor) -> None:
        """End-to-end call with HOST (ideally pinned) buffers: H2D, forward, D2H, sync."""
        with torch.cuda.device(self._device):
            check(lib().w4a16_forward_host(self._h, ctypes.c_void_p(x_host.data_ptr()),
                                           ctypes.c_void_p(y_host.data_ptr()), x_host.shape[0], _stream_ptr()))

    def forward_row_allreduce(self, x2: torch.Tensor, y: torch.Tensor, comm_ptr: Optional[int]) -> None:
        check(lib().w4a16_forward_row_allreduce(self._h, ctypes.c_void_p(x2.data_ptr()), ctypes.c_void_p(y.data_ptr()),
                                                x2.shape[0], ctypes.c_void_p(comm_ptr or 0), _stream_ptr(x2)))

    def forward_row_fused(self, x2: torch.Tensor, y: torch.Tensor, peer_ptr: Optional[int], comm_ptr: Optional[int]) -> None:
        with torch.cuda.device(self._device):
            check(lib().w4a16_forward_row_fused(self._h, ctypes.c_void_p(x2.data_ptr()), ctypes.c_void_p(y.data_ptr()),
                                                x2.shape[0], ctypes.c_void_p(peer_ptr or 0), ctypes.c_void_p(comm_ptr or 0),
                                                _stream_ptr(x2)))

    def reserve(self, max_tokens: int) -> None:
        with torch.cuda.device(self._device):
            check(lib().w4a16_reserve(self._h, max_tokens))

    # -- parity hooks -------------------------------------------------------------------
    def dequantize(self) -> torch.Tensor:
        """Dense [K, N] bf16 on the D2 grid (awq_marlin_dequant.cu:45-68, gptq_dequant.cu:58-86)."""
        w = torch.empty((self.in_features(), self.weight_rows()), dtype=torch.bfloat16, device=self._device)
        with torch.cuda.device(self._device):
            check(lib().w4a16_dequant_into(self._h, ctypes.c_void_p(w.data_ptr()), _stream_ptr(w)))
        return w

    def unpack(self):
        k, n = self.in_features(), self.weight_rows()
        g = self.weight_groups()
        q = torch.empty((k, n), dtype=torch.uint8, device=self._device)
        z = torch.empty((g, n), dtype=torch.uint8, device=self._device)
        with torch.cuda.device(self._device):
            check(lib().w4a16_unpack_into(self._h, ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(z.data_ptr()),
                                          _stream_ptr(q)))
        return q, z

    def scales(self) -> torch.Tensor:
        s = torch.empty((self.weight_groups(), self.weight_rows()), dtype=torch.bfloat16, device=self._device)
        with torch.cuda.device(self._device):
            check(lib().w4a16_scales_into(self._h, ctypes.c_void_p(s.data_ptr()), _stream_ptr(s)))
        return s

    def dequantize_d1(self, scales: torch.Tensor) -> torch.Tensor:
        """Dense [K, N] f16 on the D1 grid of awq_dequantize_cpu (awq.rs:56-166): f32(q - z) * f32(original scale),
        cast to F16.  `scales` is the checkpoint's [G, N] tensor (the resident copy is bf16-rounded)."""
        sc = scales.to(self._device).contiguous()
        if sc.dtype not in DTYPE_CODE:
            raise W4A16Error(f"scales dtype {sc.dtype} not supported")
        w = torch.empty((self.in_features(), self.weight_rows()), dtype=torch.float16, device=self._device)
        with torch.cuda.device(self._device):
            check(lib().w4a16_dequant_d1_into(self._h, ctypes.c_void_p(sc.data_ptr()), DTYPE_CODE[sc.dtype],
                                              ctypes.c_void_p(w.data_ptr()), _stream_ptr(w)))
            torch.cuda.current_stream().synchronize()   # `sc` is a local
        return w

    def row_perm(self) -> torch.Tensor:
        """Checkpoint row of every local row (w4a16_row_perm); for an act-order ROW shard: the x column order."""
        out = torch.empty((self.in_features(),), dtype=torch.int32)
        check(lib().w4a16_row_perm(self._h, ctypes.c_void_p(out.data_ptr())))
        return out

    def weight_groups(self) -> int:
        gs = self._group_size
        return 1 if gs <= 0 else self.in_features() // gs


def _make_linear(in_features, out_features, bias, group_size, fmt, device, **tp) -> B200W4A16Linear:
    return B200W4A16Linear(in_features, out_features, bias, group_size, fmt, device, **tp)


@dataclass
class MarlinConfig(QuantizationConfig):
    """marlin.rs:88-266."""
    bits: int = 4
    group_size: int = 128
    desc_act: bool = False
    is_sym: bool = True
    scalar_type: MarlinScalarType = MarlinScalarType.UINT4B8
    use_fp32_reduce: bool = True

    @staticmethod
    def gptq_int4(group_size: int) -> "MarlinConfig":
        return MarlinConfig(4, group_size, False, True, MarlinScalarType.UINT4B8)

    @staticmethod
    def awq_int4(group_size: int) -> "MarlinConfig":
        return MarlinConfig(4, group_size, False, False, MarlinScalarType.UINT4)

    def method(self):
        return QuantizationMethod.MARLIN

    def create_linear(self, in_features, out_features, bias, device):
        check_marlin_supports_shape(out_features, in_features, self.group_size)
        fmt = FORMAT_GPTQ_SYM if self.scalar_type is MarlinScalarType.UINT4B8 else FORMAT_GPTQ_ZP
        return _make_linear(in_features, out_features, bias, self.group_size, fmt, device)


@dataclass
class AwqConfig(QuantizationConfig):
    """awq.rs:170-393."""
    bits: int = 4
    group_size: int = 128
    zero_point: bool = True
    version: str = "GEMM"
    use_marlin: bool = True
    modules_to_not_convert: list = field(default_factory=lambda: ["lm_head"])

    @staticmethod
    def int4(group_size: int) -> "AwqConfig":
        return AwqConfig(4, group_size)

    @staticmethod
    def from_detected(bits, group_size, raw_config: Mapping) -> "AwqConfig":
        bits = 4 if bits is None else bits
        group_size = 128 if group_size is None else group_size
        version = "GEMV" if raw_config.get("version") in ("GEMV", "gemv") else "GEMM"
        mods = raw_config.get("modules_to_not_convert")
        mods = ["lm_head"] if not isinstance(mods, list) else [m for m in mods if isinstance(m, str)]
        return AwqConfig(bits, group_size, bool(raw_config.get("zero_point", True)), version,
                         bool(raw_config.get("is_marlin_format", False)) or bits == 4, mods)

    def method(self):
        return QuantizationMethod.AWQ

    def num_groups(self, in_features: int) -> int:
        return 1 if self.group_size <= 0 else -(-in_features // self.group_size)

    def can_use_marlin(self) -> bool:
        return self.bits == 4 and self.group_size in MARLIN_SUPPORTED_GROUP_SIZES

    def can_use_marlin_for_shape(self, in_features: int, out_features: int) -> bool:
        if not self.can_use_marlin():
            return False
        try:
            check_marlin_supports_shape(out_features, in_features, self.group_size)
            return True
        except ValueError:
            return False

    def to_marlin_config(self) -> Optional[MarlinConfig]:
        return MarlinConfig.awq_int4(self.group_size) if self.can_use_marlin() else None

    def is_layer_skipped(self, layer_name: str) -> bool:
        return any(p in layer_name for p in self.modules_to_not_convert)

    def create_linear(self, in_features, out_features, bias, device, **tp):
        if self.use_marlin and self.can_use_marlin_for_shape(in_features, out_features):
            return _make_linear(in_features, out_features, bias, self.group_size, FORMAT_AWQ, device, **tp)
        raise W4A16Error(
            f"AWQ layer {in_features}x{out_features} (bits={self.bits}, group_size={self.group_size}) is outside the "
            "Marlin gate; the reference would fall back to its CPU-dequant AwqLinear (awq.rs:384-393), "
            "the B200 path has no CPU fallback")


@dataclass
class AwqMarlinConfig(AwqConfig):
    """awq_marlin.rs:38-114: same gate, method() reports AwqMarlin."""

    def method(self):
        return QuantizationMethod.AWQ_MARLIN


@dataclass
class GptqConfig(QuantizationConfig):
    """gptq.rs:28-246."""
    bits: int = 4
    group_size: int = 128
    desc_act: bool = False
    sym: bool = True
    damp_percent: float = 0.01
    use_marlin: bool = True
    # G1 vs G3 (SURVEY 8a hazard iii): the production loader builds GptqLinear, which uses the stored
    # zero points (gptq_dequant.cu:58-86); GptqConfig::create_linear's Marlin route uses q-8.
    use_stored_zeros: bool = True

    @staticmethod
    def int4(group_size: int) -> "GptqConfig":
        return GptqConfig(4, group_size)

    @staticmethod
    def from_detected(bits, group_size, desc_act, raw_config: Mapping) -> "GptqConfig":
        bits = 4 if bits is None else bits
        return GptqConfig(bits, 128 if group_size is None else group_size, bool(desc_act),
                          bool(raw_config.get("sym", True)), float(raw_config.get("damp_percent", 0.01)),
                          bool(raw_config.get("is_marlin_format", False)) or bits == 4)

    def method(self):
        return QuantizationMethod.GPTQ

    def create_linear(self, in_features, out_features, bias, device, **tp):
        if self.bits != 4:
            raise W4A16Error(f"GPTQ bits={self.bits}: only int4 is on the B200 W4A16 path")
        check_marlin_supports_shape(out_features, in_features, self.group_size)
        fmt = FORMAT_GPTQ_ZP if self.use_stored_zeros else FORMAT_GPTQ_SYM
        return _make_linear(in_features, out_features, bias, self.group_size, fmt, device, **tp)


# ---- detection (detection.rs:240-262) ------------------------------------------------
_METHODS = {"awq": QuantizationMethod.AWQ, "awq_triton": QuantizationMethod.AWQ,
            "awq_marlin": QuantizationMethod.AWQ_MARLIN, "gptq": QuantizationMethod.GPTQ,
            "gptq_marlin": QuantizationMethod.MARLIN}


@dataclass
class DetectedQuantConfig:
    method: QuantizationMethod = QuantizationMethod.NONE
    bits: Optional[int] = None
    group_size: Optional[int] = None
    desc_act: Optional[bool] = None
    raw_config: Dict = field(default_factory=dict)


def detect_from_json(config: Mapping) -> DetectedQuantConfig:
    qc = config.get("quantization_config")
    if not isinstance(qc, Mapping):
        return DetectedQuantConfig()
    m = _METHODS.get(qc.get("quant_method"), QuantizationMethod.NONE)
    return DetectedQuantConfig(m, qc.get("bits"), qc.get("group_size"), qc.get("desc_act"), dict(qc))


def detect_from_directory(model_dir) -> DetectedQuantConfig:
    """detection.rs:70-218: config.json's `quantization_config` first, then quantize_config.json (GPTQ layout
    {bits, group_size, desc_act}; AWQ layout {w_bit, q_group_size}); anything else (missing directory, no
    quantization section, unknown method) is QuantizationMethod.NONE."""
    import json
    import os
    d = str(model_dir)
    try:
        with open(os.path.join(d, "config.json")) as f:
            det = detect_from_json(json.load(f))
        if det.method is not QuantizationMethod.NONE:
            return det
    except (OSError, ValueError):
        pass
    try:
        with open(os.path.join(d, "quantize_config.json")) as f:
            qc = json.load(f)
    except (OSError, ValueError):
        return DetectedQuantConfig()
    if not isinstance(qc, Mapping):
        return DetectedQuantConfig()
    if qc.get("bits") is not None and qc.get("group_size") is not None:   # GPTQ layout (detection.rs:180-201)
        m = QuantizationMethod.MARLIN if qc.get("quant_method") == "gptq_marlin" else QuantizationMethod.GPTQ
        return DetectedQuantConfig(m, qc.get("bits"), qc.get("group_size"), qc.get("desc_act"), dict(qc))
    if qc.get("w_bit") is not None:                                        # AWQ layout (detection.rs:203-216)
        return DetectedQuantConfig(QuantizationMethod.AWQ, qc.get("w_bit"), qc.get("q_group_size"), None, dict(qc))
    return DetectedQuantConfig()


def is_supported(capability: int, method: QuantizationMethod) -> bool:
    """quantization/mod.rs:237-264 for the W4A16 members (the reference's minimum capabilities; this build itself
    runs on capability 100 only -- QuantizationConfig.min_capability())."""
    return capability >= {QuantizationMethod.NONE: 0, QuantizationMethod.GPTQ: 70, QuantizationMethod.AWQ: 70,
                          QuantizationMethod.MARLIN: 80, QuantizationMethod.AWQ_MARLIN: 80}[method]


def create_config(detected: DetectedQuantConfig) -> QuantizationConfig:
    """quantization/mod.rs:135-213 for the W4A16 members."""
    if detected.method is QuantizationMethod.AWQ:
        return AwqConfig.from_detected(detected.bits, detected.group_size, detected.raw_config)
    if detected.method is QuantizationMethod.AWQ_MARLIN:
        c = AwqConfig.from_detected(detected.bits, detected.group_size, detected.raw_config)
        return AwqMarlinConfig(c.bits, c.group_size, c.zero_point, c.version, c.use_marlin, c.modules_to_not_convert)
    if detected.method in (QuantizationMethod.GPTQ, QuantizationMethod.MARLIN):
        return GptqConfig.from_detected(detected.bits, detected.group_size, detected.desc_act, detected.raw_config)
    raise W4A16Error(f"quantization method {detected.method} is not on the W4A16 path")


# ---- weight loaders (weight_loader.rs) --------------------------------------------------
class _Loader:
    """`tensors` plays the VarBuilder: a mapping "prefix.name" -> tensor (any device), e.g. a
    checkpoint.SafetensorsCheckpoint.  cache_dir: directory of W4T cache files (save_tiled / load_tiled): a layer
    whose cache file exists and matches is read back instead of repacked, else it is repacked and written."""

    def __init__(self, tensors: Mapping[str, torch.Tensor], config, device, cache_dir=None):
        self.tensors, self.config, self._device = tensors, config, torch.device(device)
        self.cache_dir = cache_dir
        self.cache_hits = 0

    def device(self):
        return self._device

    def dtype(self):
        return torch.bfloat16

    def _get(self, prefix, name, shape, what):
        t = self.tensors.get(f"{prefix}.{name}")
        if t is None:
            raise W4A16Error(f"{what} {name} load failed for {prefix} with expected shape {list(shape)}: missing tensor")
        if tuple(t.shape) != tuple(shape):
            raise W4A16Error(f"{what} {name} load failed for {prefix} with expected shape {list(shape)}: "
                             f"got {list(t.shape)}")
        return t

    # -- shared by the AWQ and GPTQ loaders --------------------------------------------------------------------
    def _weights(self, prefix: str, in_features: int, out_features: int, bias: bool) -> Dict[str, torch.Tensor]:
        raise NotImplementedError

    def _fingerprint(self, prefixes, weights_list) -> int:
        """Identity of the source tensors for the cache: names, dtypes, shapes and (cheaply) contents --
        the checkpoint's own hash when the mapping has one, else a hash of the packed tensors' bytes."""
        h = hashlib.blake2b(digest_size=8)
        ident = getattr(self.tensors, "fingerprint", None)
        for prefix, w in zip(prefixes, weights_list):
            for k in sorted(w):
                t = w[k]
                h.update(f"{prefix}.{k}:{t.dtype}:{list(t.shape)};".encode())
                if ident is None:
                    h.update(t.detach().cpu().contiguous().view(torch.uint8).numpy().tobytes())
        if ident is not None:
            h.update(str(ident() if callable(ident) else ident).encode())
        return int.from_bytes(h.digest(), "little") or 1

    def _cache_path(self, prefixes, tp) -> Optional[str]:
        if not self.cache_dir:
            return None
        tag = "+".join(prefixes).replace("/", "_")
        if tp:
            tag += f".tp{tp.get('tp_mode', 0)}r{tp.get('tp_rank', 0)}of{tp.get('tp_size', 1)}"
        return os.path.join(os.fspath(self.cache_dir), tag + ".w4t")

    def _load_cached(self, lin, path, fp, do_load) -> None:
        if path and os.path.exists(path):
            try:
                lin.load_tiled(path, fp)
                self.cache_hits += 1
                return
            except W4A16Error:
                pass   # stale or foreign file: repack and overwrite it
        do_load()
        if path:
            os.makedirs(os.path.dirname(path), exist_ok=True)
            lin.save_tiled(path, fp)

    def load_linear(self, prefix: str, in_features: int, out_features: int, bias: bool, **tp) -> QuantizedLinear:
        if self.config.is_layer_skipped(prefix):
            raise W4A16Error(f"{prefix} is in modules_to_not_convert: it is a dense layer, not on the W4A16 path")
        w = self._weights(prefix, in_features, out_features, bias)
        lin = self.config.create_linear(in_features, out_features, "bias" in w, self._device, **tp)
        path = self._cache_path([prefix], tp)
        fp = self._fingerprint([prefix], [w]) if path else 0
        self._load_cached(lin, path, fp, lambda: lin.load_weights(w))
        return lin

    def load_fused(self, prefixes: Sequence[str], in_features: int, out_features: Sequence[int], bias: bool,
                   epilogue: int = EPILOGUE_NONE) -> QuantizedLinear:
        """Sibling projections of one input as ONE linear (w4a16_load_*_part): e.g. prefixes
        [..q_proj, ..k_proj, ..v_proj] -> y = [q | k | v], or [..gate_proj, ..up_proj] with EPILOGUE_SWIGLU ->
        y = silu(gate) * up.  The reference creates and runs them one by one (qwen3_quantized.rs:84-96, 179-181,
        532-551).  Each sibling keeps its own checkpoint tensors; nothing is concatenated on the host."""
        if len(prefixes) != len(out_features) or not prefixes:
            raise W4A16Error("load_fused: one out_features entry per prefix")
        for prefix in prefixes:
            if self.config.is_layer_skipped(prefix):
                raise W4A16Error(f"{prefix} is in modules_to_not_convert: it is a dense layer, not on the W4A16 path")
        ws = [self._weights(pfx, in_features, n, bias) for pfx, n in zip(prefixes, out_features)]
        has_bias = ["bias" in w for w in ws]
        if any(has_bias) and not all(has_bias):
            raise W4A16Error("load_fused: either every sibling has a bias or none has")
        lin = self.config.create_linear(in_features, sum(out_features), has_bias[0], self._device)
        if epilogue != EPILOGUE_NONE:
            lin.set_epilogue(epilogue)
        path = self._cache_path(list(prefixes) + ([f"ep{epilogue}"] if epilogue else []), None)
        fp = self._fingerprint(prefixes, ws) if path else 0

        def do_load():
            n0 = 0
            for w, n in zip(ws, out_features):
                lin.load_part(n0, w)
                n0 += n
        self._load_cached(lin, path, fp, do_load)
        return lin

    def load_swiglu(self, gate_prefix: str, up_prefix: str, hidden_size: int, intermediate_size: int,
                    bias: bool = False) -> QuantizedLinear:
        """QuantizedSwiGluMlp's gate_proj + up_proj + fused_swiglu (qwen3_quantized.rs:501-551) as one linear."""
        return self.load_fused([gate_prefix, up_prefix], hidden_size, [intermediate_size, intermediate_size], bias,
                               EPILOGUE_SWIGLU)


class AwqWeightLoader(_Loader):
    """weight_loader.rs:596-714."""

    def method(self):
        return QuantizationMethod.AWQ

    def _weights(self, prefix, in_features, out_features, bias):
        packed_out = -(-out_features // 8)
        g = self.config.num_groups(in_features)
        w = {"qweight": self._get(prefix, "qweight", (in_features, packed_out), "AWQ"),
             "scales": self._get(prefix, "scales", (g, out_features), "AWQ"),
             "qzeros": self._get(prefix, "qzeros", (g, packed_out), "AWQ")}
        if bias and f"{prefix}.bias" in self.tensors:
            w["bias"] = self.tensors[f"{prefix}.bias"]
        return w


class GptqWeightLoader(_Loader):
    """weight_loader.rs:253-317."""

    def method(self):
        return QuantizationMethod.GPTQ

    def _weights(self, prefix, in_features, out_features, bias):
        gs = self.config.group_size
        g = 1 if gs <= 0 else -(-in_features // gs)
        w = {"qweight": self._get(prefix, "qweight", (in_features // 8, out_features), "GPTQ"),
             "scales": self._get(prefix, "scales", (g, out_features), "GPTQ"),
             "qzeros": self._get(prefix, "qzeros", (g, out_features // 8), "GPTQ")}
        if self.config.desc_act and f"{prefix}.g_idx" in self.tensors:  # weight_loader.rs:301-306
            w["g_idx"] = self.tensors[f"{prefix}.g_idx"]
        if bias and f"{prefix}.bias" in self.tensors:
            w["bias"] = self.tensors[f"{prefix}.bias"]
        return w


def create_weight_loader_with_params(tensors: Mapping[str, torch.Tensor], detected: DetectedQuantConfig, device,
                                     cache_dir=None):
    """weight_loader.rs:1596-1706 for the W4A16 members."""
    cfg = create_config(detected)
    if isinstance(cfg, AwqConfig):
        return AwqWeightLoader(tensors, cfg, device, cache_dir)
    return GptqWeightLoader(tensors, cfg, device, cache_dir)


class DecodeChain:
    """A fixed sequence of decode linears (M <= 4 tokens) run by one call -- the reference's per-layer call order
    (qwen3_quantized.rs:704-709 q/k/v/o, :532-551 gate/up -> silu*mul -> down) as its CUDA-graph replay sees it
    (engine/cuda_graph.rs).  `steps` is a list of (linear, x, y): x [M, in_features] and y [M, out_features] bf16
    CUDA tensors that stay alive and in place for the life of the chain (like graph arguments); linear i+1 may read
    what linear i wrote.  Where the shapes allow it run() is ONE persistent launch (`kernel()` == "chain_imma"),
    otherwise the chain calls the linears one by one."""

    def __init__(self, steps: Sequence, prologues: Optional[Sequence] = None, peers=None):
        """prologues: per step None or a dict {norm_weight, eps, residual_in, residual_out, normed_out,
        add_running_residual}: residual add + RMSNorm done by the chain kernel while it stages that linear's input
        (input_layernorm / post_attention_layernorm + the residual adds of qwen3_quantized.rs:1043-1075)."""
        steps = list(steps)
        if not steps:
            raise W4A16Error("DecodeChain: need at least one (linear, x, y) step")
        m = steps[0][1].shape[0]
        self._keep = steps
        n = len(steps)
        hs, xs, ys = (ctypes.c_void_p * n)(), (ctypes.c_void_p * n)(), (ctypes.c_void_p * n)()
        for i, (lin, x, y) in enumerate(steps):
            inner = getattr(lin, "inner", lin)   # Column/RowParallel wrappers expose their local linear
            inner._check_x(x)
            if x.dim() != 2 or y.dim() != 2 or not x.is_contiguous() or not y.is_contiguous():
                raise W4A16Error("DecodeChain: x and y must be contiguous rank-2 tensors")
            if x.shape[0] != m or y.shape[0] != m:
                raise W4A16Error(f"DecodeChain: every step must have the same M = {m}")
            if y.dtype != torch.bfloat16 or y.shape[1] != inner.out_features() or not y.is_cuda:
                raise W4A16Error(f"DecodeChain: y of step {i} must be a CUDA bf16 tensor [M, {inner.out_features()}]")
            hs[i], xs[i], ys[i] = inner._h, x.data_ptr(), y.data_ptr()
        self._device = steps[0][1].device
        self._c = ctypes.c_void_p()
        pro = None
        if prologues is not None:
            from ._lib import ChainPrologue
            if len(prologues) != n:
                raise W4A16Error("DecodeChain: prologues must have one entry (or None) per step")
            pro = (ChainPrologue * n)()
            self._keep_pro = prologues
            for i, pr in enumerate(prologues):
                if pr is None:
                    continue
                k = steps[i][1].shape[1]
                w = pr["norm_weight"]
                if w.dtype != torch.bfloat16 or w.numel() != k or not w.is_cuda or not w.is_contiguous():
                    raise W4A16Error(f"DecodeChain: norm_weight of step {i} must be a contiguous CUDA bf16 tensor [{k}]")
                pro[i].norm_weight = w.data_ptr()
                for key in ("residual_in", "residual_out", "normed_out"):
                    t = pr.get(key)
                    if t is not None:
                        if t.dtype != torch.bfloat16 or tuple(t.shape) != (m, k) or not t.is_cuda or not t.is_contiguous():
                            raise W4A16Error(f"DecodeChain: {key} of step {i} must be a contiguous CUDA bf16 tensor [{m}, {k}]")
                        setattr(pro[i], key, t.data_ptr())
                pro[i].eps = float(pr.get("eps", 1e-6))
                pro[i].add_running_residual = int(bool(pr.get("add_running_residual", False)))
        check(lib().w4a16_chain_create_ex(ctypes.byref(self._c), n, hs, xs, ys, m, pro))
        if peers is not None:   # tensor-parallel: the ROW shards all-reduce inside the launch (parallel.PeerGroup)
            self._peers = peers
            check(lib().w4a16_chain_set_peer(self._c, peers.ptr()))
        self._n, self._m = n, m

    def __del__(self):
        c = getattr(self, "_c", None)
        if c:
            try:
                lib().w4a16_chain_destroy(c)
            except Exception:
                pass
            self._c = None

    def __len__(self) -> int:
        return self._n

    def kernel(self) -> str:
        return lib().w4a16_chain_kernel(self._c).decode()

    def run(self) -> None:
        check(lib().w4a16_chain_run(self._c, torch.cuda.current_stream(self._device).cuda_stream))

    def run_host(self, x0_host: torch.Tensor, y_last_host: torch.Tensor) -> None:
        """run() with HOST (ideally pinned) buffers at the two ends of the chain: H2D, chain, D2H, sync."""
        with torch.cuda.device(self._device):
            check(lib().w4a16_chain_run_host(self._c, ctypes.c_void_p(x0_host.data_ptr()),
                                             ctypes.c_void_p(y_last_host.data_ptr()),
                                             torch.cuda.current_stream(self._device).cuda_stream))

    def debug_stamps(self):
        """[n, CTAs, 8] %globaltimer stamps of the last run (W4A16_GV_DEBUG=1 at creation)."""
        import numpy as np
        sms = torch.cuda.get_device_properties(self._device).multi_processor_count
        buf = np.zeros((self._n, sms, 8), dtype=np.uint64)
        check(lib().w4a16_chain_debug_stamps(self._c, ctypes.c_void_p(buf.ctypes.data), buf.size))
        return buf


def kernel_launches() -> int:
    return int(lib().w4a16_kernel_launches())


def set_dispatch(mode: str) -> None:
    check(lib().w4a16_set_dispatch(mode.encode()))
