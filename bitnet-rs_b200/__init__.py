"""bitnet-rs_b200 -- B200 (sm_100a) implementation of BitNet-rs's I2_S ternary linear layer.

This package is the Python face of libbitnet_b200.so (C ABI in include/bitnet_b200.h). It
mirrors the reference's kernel interface -- `KernelProvider`, `KernelManager`,
`QuantizationType`, `gemv_qk256`, `I2SQk256NoScale` (crates/bitnet-kernels/src/lib.rs:39-176,
crates/bitnet-quantization/src/i2s_qk256.rs:65-128,346-372) -- and adds the device-resident
fast path. There is no CPU fallback: without the compiled CUDA library every call raises.

Import as `bitnet_rs_b200` (the shim module at the repo root maps the hyphenated directory).
"""
from .ffi import LIB_PATH, B200Error, build, lib, last_error, launch_count, last_linear_kernel  # noqa: F401
from .provider import (  # noqa: F401
    B200Kernel,
    I2SQk256NoScale,
    KernelError,
    KernelManager,
    KernelProvider,
    QuantizationType,
    QuantizedLinear,
    gemv_qk256,
    i2s_matmul_f32,
    qk256_gemv,
    qk256_gemv_scalar,
    quantize_input_i2s,
    unpack_2bit_values,
)
from .device import (  # noqa: F401
    EPI_RESIDUAL,
    EPI_STORE,
    EPI_SWIGLU,
    LAYOUT_BITNET32_F16,
    LAYOUT_QK256,
    MAP_A,
    MAP_B,
    MAP_C,
    SCALE_ABS_CLAMP,
    SCALE_RAW,
    Context,
    Weight,
)

from .model import (  # noqa: F401
    Model,
    ModelShape,
    rope_tables,
    synth_bytes,
    synth_embedding,
    synth_layer,
    synth_uniform,
    synth_weight,
)

from . import gguf, tp  # noqa: F401,E402
from .tp import TPDecoder, TpFabric  # noqa: F401,E402

__all__ = [
    "TPDecoder", "TpFabric", "tp",
    "Model", "ModelShape",
    "B200Error", "B200Kernel", "Context", "I2SQk256NoScale", "KernelError", "KernelManager",
    "KernelProvider", "QuantizationType", "Weight", "build", "gemv_qk256", "i2s_matmul_f32",
    "lib", "LAYOUT_QK256", "LAYOUT_BITNET32_F16", "MAP_A", "MAP_B", "MAP_C",
]
