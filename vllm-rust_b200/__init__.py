"""b200-w4a16: host-side mirror of vllm-rust's quantized-linear API
(crates/core/src/quantization/{config,awq,awq_marlin,marlin,gptq,weight_loader,detection}.rs)
over the C ABI of libw4a16_b200.so.  Python stands in for the Rust shim because
this image has no Rust toolchain; INTEGRATION.md holds the Rust binding."""
from ._lib import W4A16Error, build, lib, LIB_PATH  # noqa: F401
from .quantization import (  # noqa: F401
    MARLIN_SUPPORTED_GROUP_SIZES, GPTQ_MARLIN_MIN_THREAD_N, GPTQ_MARLIN_MIN_THREAD_K,
    QuantizationMethod, QuantizationConfig, QuantizedLinear,
    AwqConfig, AwqMarlinConfig, GptqConfig, MarlinConfig, MarlinScalarType,
    B200W4A16Linear, check_marlin_supports_shape,
    AwqWeightLoader, GptqWeightLoader, create_weight_loader_with_params,
    detect_from_json, detect_from_directory, is_supported, create_config, kernel_launches, set_dispatch,
    DetectedQuantConfig, EPILOGUE_NONE, EPILOGUE_SWIGLU, DecodeChain,
)
from .checkpoint import SafetensorsCheckpoint, load_decoder_linears, save_safetensors, save_sharded  # noqa: F401
from .parallel import (  # noqa: F401
    DeviceCommunicator, MockCommunicator, NcclCommunicator, PeerGroup, TorchDistCommunicator,
    ColumnParallelW4A16Linear, RowParallelW4A16Linear, shard_range,
)
