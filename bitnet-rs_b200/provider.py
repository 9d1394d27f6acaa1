"""Host-side mirror of the reference's kernel interface, over the C ABI.

Same names, argument meaning and error behaviour as
  trait KernelProvider / KernelManager   crates/bitnet-kernels/src/lib.rs:39-176
  FfiKernel                              crates/bitnet-kernels/src/ffi/bridge.rs:159-263
  gemv_qk256 / I2SQk256NoScale           crates/bitnet-quantization/src/i2s_qk256.rs:65-128,346-372
  QuantizationType                       crates/bitnet-common/src/types.rs:7-14
  KernelError                            crates/bitnet-common/src/error.rs:80-97
so the parity tests read like the reference's own. Slices are numpy arrays (host memory);
every call is synchronous, as the reference's providers are.
"""
import enum

import numpy as np

from . import ffi


class QuantizationType(enum.Enum):
    I2S = 0
    TL1 = 1
    TL2 = 2

    def __str__(self):
        return {0: "I2_S", 1: "TL1", 2: "TL2"}[self.value]


class KernelError(Exception):
    """bitnet_common::KernelError; `kind` is the variant name."""

    def __init__(self, kind, reason=""):
        self.kind, self.reason = kind, reason
        text = {
            "NoProvider": "No available kernel provider",
            "ExecutionFailed": "Kernel execution failed: %s" % reason,
            "GpuError": "GPU error: %s" % reason,
            "InvalidArguments": "Invalid arguments: %s" % reason,
        }.get(kind, "%s: %s" % (kind, reason))
        super().__init__(text)


class KernelProvider:
    """trait KernelProvider (lib.rs:39-58)."""

    def name(self):
        raise NotImplementedError

    def is_available(self):
        raise NotImplementedError

    def matmul_i2s(self, a, b, c, m, n, k):
        raise NotImplementedError

    def quantize(self, input, output, scales, qtype):
        raise NotImplementedError


def _host(arr, dtype, what):
    if not isinstance(arr, np.ndarray) or arr.dtype != dtype or not arr.flags["C_CONTIGUOUS"]:
        raise TypeError("%s must be a C-contiguous numpy array of %s" % (what, np.dtype(dtype).name))
    return arr


class B200Kernel(KernelProvider):
    """The drop-in provider: FfiKernel's shape (bridge.rs:159-263) bound to libbitnet_b200.so.

    `new` calls bitnet_cpp_init, `close`/GC calls bitnet_cpp_cleanup; several instances may
    coexist (the C side reference-counts)."""

    def __init__(self):
        self._lib = ffi.lib()
        self.initialized = self._lib.bitnet_cpp_init() == 0
        if not self.initialized:
            raise KernelError("ExecutionFailed", ffi.last_error() or "BitNet B200 kernels not available")

    def close(self):
        if getattr(self, "initialized", False):
            self._lib.bitnet_cpp_cleanup()
            self.initialized = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def name(self):
        return "b200"

    def is_available(self):
        return self.initialized and self._lib.bitnet_cpp_is_available() != 0

    def matmul_i2s(self, a, b, c, m, n, k):
        if not self.is_available():
            raise KernelError("NoProvider")
        a = _host(a, np.int8, "a")
        b = _host(b, np.uint8, "b")
        c = _host(c, np.float32, "c")
        # same validation, same messages as bridge.rs:205-220 / fallback.rs:48-63
        if a.size != m * k:
            raise KernelError("ExecutionFailed",
                              "Matrix A dimension mismatch: expected %d, got %d" % (m * k, a.size))
        if b.size != k * n:
            raise KernelError("ExecutionFailed",
                              "Matrix B dimension mismatch: expected %d, got %d" % (k * n, b.size))
        if c.size != m * n:
            raise KernelError("ExecutionFailed",
                              "Matrix C dimension mismatch: expected %d, got %d" % (m * n, c.size))
        if m * n == 0:
            return
        if k == 0:
            c.fill(0.0)
            return
        rc = self._lib.bitnet_cpp_matmul_i2s(a.ctypes.data, b.ctypes.data, c.ctypes.data, m, n, k)
        if rc != 0:
            raise KernelError("ExecutionFailed", "C++ kernel error: %s" % ffi.last_error())

    def quantize(self, input, output, scales, qtype):
        if not self.is_available():
            raise KernelError("NoProvider")
        input = _host(input, np.float32, "input")
        output = _host(output, np.uint8, "output")
        scales = _host(scales, np.float32, "scales")
        qtype = QuantizationType(qtype) if not isinstance(qtype, QuantizationType) else qtype
        rc = self._lib.bitnet_cpp_quantize(input.ctypes.data, input.size, output.ctypes.data,
                                           output.size, scales.ctypes.data, scales.size, qtype.value)
        if rc != 0:
            raise KernelError("ExecutionFailed", "C++ quantization error: %s" % ffi.last_error())


class KernelManager:
    """KernelManager (lib.rs:61-176) with the B200 provider as its only entry: north_star
    forbids multi-backend dispatch and CPU fallback, so an unavailable device is NoProvider."""

    def __init__(self):
        self.providers = []
        try:
            self.providers.append(B200Kernel())
        except (KernelError, ffi.B200Error):
            pass
        self._selected = None

    def select_best(self):
        if self._selected is None:
            for i, p in enumerate(self.providers):
                if p.is_available():
                    self._selected = i
                    break
        if self._selected is None:
            raise KernelError("NoProvider")
        return self.providers[self._selected]

    def selected_provider_name(self):
        return None if self._selected is None else self.providers[self._selected].name()

    def list_available_providers(self):
        return [p.name() for p in self.providers if p.is_available()]


QK256_BLOCK = 256
QK256_PACKED_BYTES = 64


class I2SQk256NoScale:
    """I2SQk256NoScale (i2s_qk256.rs:65-128): rows, cols, row_stride_bytes, qs."""

    TOLERANCE = 128

    def __init__(self, rows, cols, qs):
        qs = np.ascontiguousarray(qs, dtype=np.uint8).ravel()
        blocks_per_row = -(-cols // QK256_BLOCK)
        self.row_stride_bytes = blocks_per_row * QK256_PACKED_BYTES
        expected = rows * self.row_stride_bytes
        if abs(qs.size - expected) > self.TOLERANCE:
            raise ValueError(
                "I2SQk256NoScale: data size mismatch: got %d bytes, expected %d for %dx%d matrix. "
                "Check tensor orientation: QK256 requires [out_dim, in_dim] layout."
                % (qs.size, expected, rows, cols))
        self.rows, self.cols, self.qs = rows, cols, qs

    def row_bytes(self, row):
        s = row * self.row_stride_bytes
        return self.qs[s:s + self.row_stride_bytes]


def gemv_qk256(qs_data, x, y_out, rows, cols, row_stride_bytes):
    """gemv_qk256 (i2s_qk256.rs:346-372): y = A x, A packed QK256, x f32, on host slices.

    Raises ValueError with the reference's messages ("y_out length", "x length", "too short")."""
    qs_data = _host(qs_data, np.uint8, "qs_data")
    x = _host(x, np.float32, "x")
    y_out = _host(y_out, np.float32, "y_out")
    rc = ffi.lib().bitnet_b200_gemv_qk256(qs_data.ctypes.data, qs_data.size, x.ctypes.data, x.size,
                                          y_out.ctypes.data, y_out.size, rows, cols, row_stride_bytes)
    if rc != 0:
        raise ValueError(ffi.last_error())


def i2s_matmul_f32(activations, weights_packed, scales, out, m, n, k, block_size):
    """launch_i2s_matmul (cuda/quantized_matmul.rs:282-303) with the numerics of
    i2s_matmul_f32 (cpu/quantized_matmul.rs:57-96): map C, packed [n, ceil(k/4)], f32 scales."""
    activations = _host(activations, np.float32, "activations")
    weights_packed = _host(weights_packed, np.uint8, "weights_packed")
    scales = _host(scales, np.float32, "scales")
    out = _host(out, np.float32, "out")
    # validate_matmul_args, cpu/quantized_matmul.rs:204-255
    if block_size == 0:
        raise KernelError("ExecutionFailed", "block_size must be > 0")
    if m == 0 or n == 0 or k == 0:
        raise KernelError("ExecutionFailed", "dimensions must be > 0: m=%d, n=%d, k=%d" % (m, n, k))
    packed_k, nbk = -(-k // 4), -(-k // block_size)
    if activations.size < m * k:
        raise KernelError("ExecutionFailed",
                          "activations too small: expected %d, got %d" % (m * k, activations.size))
    if weights_packed.size < packed_k * n:
        raise KernelError("ExecutionFailed", "weights_packed too small: expected %d, got %d"
                          % (packed_k * n, weights_packed.size))
    if scales.size < n * nbk:
        raise KernelError("ExecutionFailed", "scales too small: expected %d, got %d" % (n * nbk, scales.size))
    if out.size < m * n:
        raise KernelError("ExecutionFailed", "output too small: expected %d, got %d" % (m * n, out.size))
    rc = ffi.lib().bitnet_b200_i2s_matmul_f32(activations.ctypes.data, weights_packed.ctypes.data,
                                              scales.ctypes.data, out.ctypes.data, m, n, k, block_size)
    if rc != 0:
        raise KernelError("ExecutionFailed", ffi.last_error())


# --------------------------------------------------------------------------------------------
# a12: the legacy QK256 entry points (crates/bitnet-quantization/src/qk256_dispatch.rs:20-99)
# --------------------------------------------------------------------------------------------
def _legacy(fn, output, rows, cols, packed, scales, activations):
    output = _host(output, np.float32, "output")
    packed = _host(packed, np.uint8, "packed")
    scales = _host(scales, np.float32, "scales")
    activations = _host(activations, np.float32, "activations")
    rc = fn(output.ctypes.data, output.size, rows, cols, packed.ctypes.data, packed.size, scales.ctypes.data,
            scales.size, activations.ctypes.data, activations.size)
    if rc != 0:
        # the reference asserts (panics) on a violated contract: same message, as an AssertionError
        raise AssertionError(ffi.last_error())


def qk256_gemv(output, rows, cols, packed, scales, activations):
    """qk256_gemv (qk256_dispatch.rs:20-43): validates the legacy contract, then gemv_qk256 (map A);
    `scales` is checked for its length and not consumed."""
    _legacy(ffi.lib().bitnet_b200_qk256_gemv, output, rows, cols, packed, scales, activations)


def qk256_gemv_scalar(output, rows, cols, packed, scales, activations):
    """qk256_gemv_scalar (qk256_dispatch.rs:45-99): map D {-1, 0, +1, -1} with one f32 scale per 256-block."""
    _legacy(ffi.lib().bitnet_b200_qk256_gemv_scalar, output, rows, cols, packed, scales, activations)


# --------------------------------------------------------------------------------------------
# a16: the in-tree consumer of the provider seam, QuantizedLinear's native I2_S path
# (crates/bitnet-inference/src/layers/quantized_linear.rs:704-802, utils at :1762-1773 and
# crates/bitnet-quantization/src/utils.rs:76-91). Host-side glue only; the matmul is the provider's.
# --------------------------------------------------------------------------------------------
def unpack_2bit_values(packed, output_len):
    """unpack_2bit_values (bitnet-quantization/src/utils.rs:76-91): 4 codes per byte, LSB first, code - 2."""
    packed = np.ascontiguousarray(packed, dtype=np.uint8).reshape(-1)
    codes = ((packed[:, None] >> (2 * np.arange(4, dtype=np.uint8))) & 3).reshape(-1)[:output_len]
    return codes.astype(np.int8) - 2


def quantize_input_i2s(x):
    """quantize_input_i2s (quantized_linear.rs:1762-1773): clamp to [-2, 1], round half away from zero (f32::round)."""
    c = np.clip(np.asarray(x, dtype=np.float32), -2.0, 1.0)
    return (np.sign(c) * np.floor(np.abs(c) + np.float32(0.5))).astype(np.int8)


class QuantizedLinear:
    """QuantizedLinear's I2_S branch: forward = quantize_input_i2s -> provider.matmul_i2s on the unpacked
    (code + 2) bytes -> apply_quantization_scales. `weights_packed` holds in*out 2-bit codes of the [in, out]
    (row-major k x n) matrix matmul_i2s expects; `scales` are per output feature or per `block_size` weights."""

    def __init__(self, weights_packed, scales, in_features, out_features, block_size=32, provider=None):
        self.data = np.ascontiguousarray(weights_packed, dtype=np.uint8).reshape(-1)
        self.scales = np.ascontiguousarray(scales, dtype=np.float32).reshape(-1)
        self.in_features, self.out_features, self.block_size = int(in_features), int(out_features), int(block_size)
        self.provider = provider if provider is not None else B200Kernel()

    def prepare_quantized_weights_i2s(self):
        """quantized_linear.rs:769-776: unpack to i8 in [-2, 1], shift to u8 in [0, 3] for the kernel interface."""
        return (unpack_2bit_values(self.data, self.in_features * self.out_features) + 2).astype(np.uint8)

    def apply_quantization_scales(self, output_data):
        """quantized_linear.rs:779-802 (in place on [batch, out_features] f32): scales[col] when there is one scale
        per output feature, else scales[min(col * in_features / block_size, len - 1)]; a missing scale counts as 1."""
        out = self.out_features
        n = self.scales.size
        if n == out:
            idx = np.arange(out)
        else:
            idx = np.minimum(np.arange(out, dtype=np.int64) * self.in_features // self.block_size, n - 1)
        sc = self.scales[idx] if n else np.ones(out, np.float32)
        output_data.reshape(-1, out)[:] *= np.float32(1.0) * sc
        return output_data

    def forward(self, x):
        x = np.asarray(x, dtype=np.float32)
        batch = int(np.prod(x.shape[:-1])) if x.ndim > 1 else 1
        q = quantize_input_i2s(x.reshape(batch, self.in_features))
        w = self.prepare_quantized_weights_i2s()
        y = np.zeros(batch * self.out_features, np.float32)
        self.provider.matmul_i2s(q.reshape(-1), w, y, batch, self.out_features, self.in_features)
        self.apply_quantization_scales(y)
        return y.reshape(x.shape[:-1] + (self.out_features,))
