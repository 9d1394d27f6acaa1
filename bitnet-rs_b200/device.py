"""Device-resident fast path (Part 3 of include/bitnet_b200.h) on torch CUDA tensors.

torch is plumbing only: it owns device memory and the stream; every kernel that runs is
from libbitnet_b200.so. Nothing here falls back to torch ops.
"""
import ctypes as C

import numpy as np

from . import ffi

LAYOUT_QK256 = 0
LAYOUT_BITNET32_F16 = 1
SCALE_ABS_CLAMP = 0
SCALE_RAW = 1
EPI_STORE = 0
EPI_RESIDUAL = 1
EPI_SWIGLU = 2
MAP_A = (-2, -1, 1, 2)   # i2s_qk256.rs:139-146 -- production map of both layouts
MAP_B = (-2, -1, 0, 1)   # quantization/src/utils.rs:76-91
MAP_C = (0, 1, 0, -1)    # kernels/src/cpu/quantized_matmul.rs:19-27


def _torch():
    import torch
    return torch


def _dev_ptr(t, dtype, what):
    torch = _torch()
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == dtype and t.is_contiguous()):
        raise TypeError("%s must be a contiguous CUDA tensor of %s" % (what, dtype))
    return t.data_ptr()


class Context:
    """bitnet_b200_ctx: a device ordinal and the stream kernels are launched on.

    By default it launches on torch's current stream of that device, so torch-side
    allocation, events and CUDA-graph capture see the kernels."""

    def __init__(self, device=0, stream="torch"):
        torch = _torch()
        self._lib = ffi.lib()
        self.device = int(device)
        if stream == "torch":
            stream = torch.cuda.current_stream(self.device).cuda_stream
        h = C.c_void_p()
        ffi.check(self._lib.bitnet_b200_ctx_create(self.device, C.c_void_p(stream or None), C.byref(h)))
        self._h = h

    def set_stream(self, cuda_stream):
        """Launch on this cudaStream_t from now on (0 / None = the legacy default stream)."""
        ffi.check(self._lib.bitnet_b200_ctx_set_stream(self._h, C.c_void_p(cuda_stream or None)))

    def use_torch_stream(self):
        """Re-bind to torch's current stream (call inside `torch.cuda.graph(...)` / `stream(...)`)."""
        self.set_stream(_torch().cuda.current_stream(self.device).cuda_stream)

    def synchronize(self):
        ffi.check(self._lib.bitnet_b200_ctx_synchronize(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.bitnet_b200_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- weights ---------------------------------------------------------------------
    def weight(self, host_bytes, layout, rows, cols, lut=None, scale_policy=SCALE_ABS_CLAMP):
        return Weight(self, host_bytes, layout, rows, cols, lut, scale_policy)

    # ---- activation quantisers -------------------------------------------------------
    def actquant_i8(self, x, q=None, scale=None):
        """quantize_symmetric_i8 per token row (cpu/quantize.rs:32-49) on device."""
        torch = _torch()
        M, K = x.shape
        q = torch.empty((M, K), dtype=torch.int8, device=x.device) if q is None else q
        scale = torch.empty((M,), dtype=torch.float32, device=x.device) if scale is None else scale
        ffi.check(self._lib.bitnet_b200_actquant_i8(
            self._h, _dev_ptr(x, torch.float32, "x"), M, K, _dev_ptr(q, torch.int8, "q"),
            _dev_ptr(scale, torch.float32, "scale")))
        return q, scale

    def rmsnorm_actquant(self, x, gamma, eps=1e-5, q=None, scale=None, normed=None, want_normed=False):
        """cpu::rms_norm (cpu/layer_norm.rs:118-137) with the act-quant fused in its epilogue."""
        torch = _torch()
        M, K = x.shape
        q = torch.empty((M, K), dtype=torch.int8, device=x.device) if q is None else q
        scale = torch.empty((M,), dtype=torch.float32, device=x.device) if scale is None else scale
        if normed is None and want_normed:
            normed = torch.empty((M, K), dtype=torch.float32, device=x.device)
        ffi.check(self._lib.bitnet_b200_rmsnorm_actquant(
            self._h, _dev_ptr(x, torch.float32, "x"), _dev_ptr(gamma, torch.float32, "gamma"),
            C.c_float(eps), M, K, _dev_ptr(q, torch.int8, "q"), _dev_ptr(scale, torch.float32, "scale"),
            None if normed is None else _dev_ptr(normed, torch.float32, "normed")))
        return (q, scale, normed) if (want_normed or normed is not None) else (q, scale)

    # ---- the ternary linear ------------------------------------------------------------
    def linear_i8(self, w, q, act_scale, y=None, want_acc=False, acc_out=None):
        """acc_out: write the int32 accumulators into this tensor (e.g. a peer-visible exchange buffer)."""
        torch = _torch()
        M = q.shape[0]
        if q.shape[1] != w.cols:
            raise ValueError("q has %d columns, weight has %d" % (q.shape[1], w.cols))
        y = torch.empty((M, w.rows), dtype=torch.float32, device=q.device) if y is None else y
        acc = acc_out
        if acc_out is not None:
            if w.layout != LAYOUT_QK256 or acc_out.dtype != torch.int32 or acc_out.numel() != M * w.rows or not acc_out.is_contiguous():
                raise ValueError("acc_out must be a contiguous int32 [M, rows] tensor of a QK256 weight")
        elif want_acc:
            shape = (M, w.rows) if w.layout == LAYOUT_QK256 else (M, w.rows, w.cols // 32)
            # QK256: the kernel writes every accumulator; BitNet32-F16 blocks are zero-initialised
            acc = (torch.empty if w.layout == LAYOUT_QK256 else torch.zeros)(shape, dtype=torch.int32, device=q.device)
        ffi.check(self._lib.bitnet_b200_linear_i8(
            self._h, w._h, _dev_ptr(q, torch.int8, "q"), _dev_ptr(act_scale, torch.float32, "act_scale"),
            M, _dev_ptr(y, torch.float32, "y"), None if acc is None else acc.data_ptr()))
        return (y, acc) if (want_acc or acc_out is not None) else y

    def linear_i8_static_scratch(self, w, q, act_scale, y):
        """linear_i8 (tcgen05 GEMM, QK256, M >= 9) with the split-K scratch in the library's static device array: the call
        allocates nothing (y is the caller's). Calls that overlap on one GPU race on the shared scratch."""
        torch = _torch()
        M = q.shape[0]
        if q.shape[1] != w.cols:
            raise ValueError("q has %d columns, weight has %d" % (q.shape[1], w.cols))
        if y.shape != (M, w.rows):
            raise ValueError("y must be [M, rows]")
        ffi.check(self._lib.bitnet_b200_linear_i8_static_scratch(
            self._h, w._h, _dev_ptr(q, torch.int8, "q"), _dev_ptr(act_scale, torch.float32, "act_scale"), M,
            _dev_ptr(y, torch.float32, "y")))
        return y

    def linear_i8_swiglu(self, w, q, act_scale, h=None):
        """linear_i8 on an interleaved (gate, up) weight with SwiGLU in the GEMM epilogue: h [M, rows/2] (M >= 9)."""
        torch = _torch()
        M = q.shape[0]
        if q.shape[1] != w.cols:
            raise ValueError("q has %d columns, weight has %d" % (q.shape[1], w.cols))
        h = torch.empty((M, w.rows // 2), dtype=torch.float32, device=q.device) if h is None else h
        ffi.check(self._lib.bitnet_b200_linear_i8_swiglu(
            self._h, w._h, _dev_ptr(q, torch.int8, "q"), _dev_ptr(act_scale, torch.float32, "act_scale"), M,
            _dev_ptr(h, torch.float32, "h")))
        return h

    def fused_linear(self, w, x, gamma=None, eps=1e-5, epilogue=EPI_STORE, y=None, residual=None, want_q=False):
        """[RMSNorm ->] act-quant -> ternary linear -> epilogue in ONE launch (decode, M <= 4)."""
        torch = _torch()
        M = x.shape[0]
        if x.shape[1] != w.cols:
            raise ValueError("x has %d columns, weight has %d" % (x.shape[1], w.cols))
        n_out = w.rows // 2 if epilogue == EPI_SWIGLU else w.rows
        y = torch.empty((M, n_out), dtype=torch.float32, device=x.device) if y is None else y
        q = s = None
        if want_q:
            q = torch.empty((M, w.cols), dtype=torch.int8, device=x.device)
            s = torch.empty((M,), dtype=torch.float32, device=x.device)
        ffi.check(self._lib.bitnet_b200_fused_linear(
            self._h, w._h, _dev_ptr(x, torch.float32, "x"),
            None if gamma is None else _dev_ptr(gamma, torch.float32, "gamma"), C.c_float(eps), M, int(epilogue),
            _dev_ptr(y, torch.float32, "y"), None if residual is None else _dev_ptr(residual, torch.float32, "residual"),
            None if q is None else q.data_ptr(), None if s is None else s.data_ptr()))
        return (y, q, s) if want_q else y

    def linear_f32(self, w, x, y=None):
        torch = _torch()
        M = x.shape[0]
        y = torch.empty((M, w.rows), dtype=torch.float32, device=x.device) if y is None else y
        ffi.check(self._lib.bitnet_b200_linear_f32(
            self._h, w._h, _dev_ptr(x, torch.float32, "x"), M, _dev_ptr(y, torch.float32, "y")))
        return y

    # ---- tensor-parallel building blocks (Part 5) --------------------------------------
    def row_absmax(self, x, swiglu=False, absmax=None):
        """Per-token max|v| of the local slice; swiglu: x is [M,2K] interleaved (gate, up), returns (h, absmax)."""
        torch = _torch()
        M = x.shape[0]
        K = x.shape[1] // 2 if swiglu else x.shape[1]
        am = torch.empty((M,), dtype=torch.float32, device=x.device) if absmax is None else absmax
        h = torch.empty((M, K), dtype=torch.float32, device=x.device) if swiglu else None
        ffi.check(self._lib.bitnet_b200_row_absmax(
            self._h, _dev_ptr(x, torch.float32, "x"), M, K, 1 if swiglu else 0,
            None if h is None else h.data_ptr(), am.data_ptr()))
        return (h, am) if swiglu else am

    def actquant_i8_absmax(self, x, absmax, q=None, scale=None):
        """quantize_symmetric_i8 per row with the given (all-reduced) absmax."""
        torch = _torch()
        M, K = x.shape
        q = torch.empty((M, K), dtype=torch.int8, device=x.device) if q is None else q
        scale = torch.empty((M,), dtype=torch.float32, device=x.device) if scale is None else scale
        ffi.check(self._lib.bitnet_b200_actquant_i8_absmax(
            self._h, _dev_ptr(x, torch.float32, "x"), _dev_ptr(absmax, torch.float32, "absmax"), M, K,
            _dev_ptr(q, torch.int8, "q"), _dev_ptr(scale, torch.float32, "scale")))
        return q, scale

    def tp_xchg_max_actquant(self, xc, x, q=None, scale=None):
        """all-reduce(MAX) of the per-token absmax pulled from the peers + quantise the local slice, one kernel."""
        torch = _torch()
        M, K = x.shape
        q = torch.empty((M, K), dtype=torch.int8, device=x.device) if q is None else q
        scale = torch.empty((M,), dtype=torch.float32, device=x.device) if scale is None else scale
        ffi.check(self._lib.bitnet_b200_tp_xchg_max_actquant(
            self._h, C.byref(xc), _dev_ptr(x, torch.float32, "x"), M, K, _dev_ptr(q, torch.int8, "q"),
            _dev_ptr(scale, torch.float32, "scale")))
        return q, scale

    def tp_xchg_sum_rmsnorm_actquant(self, xc, act_scale, x, gamma, eps, q=None, scale=None, want_normed=False):
        """all-reduce(SUM, int32) pulled from the peers + residual into x + RMSNorm + act-quant, one kernel."""
        torch = _torch()
        M, N = x.shape
        q = torch.empty((M, N), dtype=torch.int8, device=x.device) if q is None else q
        scale = torch.empty((M,), dtype=torch.float32, device=x.device) if scale is None else scale
        normed = torch.empty((M, N), dtype=torch.float32, device=x.device) if want_normed else None
        ffi.check(self._lib.bitnet_b200_tp_xchg_sum_rmsnorm_actquant(
            self._h, C.byref(xc), _dev_ptr(act_scale, torch.float32, "act_scale"), _dev_ptr(x, torch.float32, "x"),
            _dev_ptr(gamma, torch.float32, "gamma"), float(eps), M, N, _dev_ptr(q, torch.int8, "q"),
            _dev_ptr(scale, torch.float32, "scale"), None if normed is None else normed.data_ptr()))
        return (q, scale, normed) if want_normed else (q, scale)

    def residual_acc(self, acc, act_scale, x):
        """x += act_scale[m] * float(acc[m, n]) in place."""
        torch = _torch()
        M, N = x.shape
        ffi.check(self._lib.bitnet_b200_residual_acc(
            self._h, _dev_ptr(acc, torch.int32, "acc"), _dev_ptr(act_scale, torch.float32, "act_scale"),
            _dev_ptr(x, torch.float32, "x"), M, N))
        return x

    def rope_attn_decode(self, qkv, cos_t, sin_t, kc, vc, pos, n_heads, n_kv_heads, head_dim, max_ctx, att=None):
        """RoPE + KV append + GQA decode attention on the local heads; pos is a device int32 scalar tensor."""
        torch = _torch()
        B = qkv.shape[0]
        att = torch.empty((B, n_heads * head_dim), dtype=torch.float32, device=qkv.device) if att is None else att
        ffi.check(self._lib.bitnet_b200_rope_attn_decode(
            self._h, _dev_ptr(qkv, torch.float32, "qkv"), _dev_ptr(cos_t, torch.float32, "cos"),
            _dev_ptr(sin_t, torch.float32, "sin"), _dev_ptr(kc, torch.float32, "kc"),
            _dev_ptr(vc, torch.float32, "vc"), _dev_ptr(pos, torch.int32, "pos"), B, n_heads, n_kv_heads,
            head_dim, max_ctx, _dev_ptr(att, torch.float32, "att")))
        return att


class Weight:
    """bitnet_b200_weight: one packed tensor uploaded once, in the reference's byte layout."""

    def __init__(self, ctx, host_bytes, layout, rows, cols, lut=None, scale_policy=SCALE_ABS_CLAMP):
        self._lib = ffi.lib()
        host = np.ascontiguousarray(host_bytes, dtype=np.uint8).ravel()
        lut_arr = None if lut is None else np.asarray(lut, dtype=np.int8)
        h = C.c_void_p()
        ffi.check(self._lib.bitnet_b200_weight_create(
            ctx._h, host.ctypes.data, host.size, int(layout), int(rows), int(cols),
            None if lut_arr is None else lut_arr.ctypes.data, int(scale_policy), C.byref(h)))
        self._h = h
        self.layout, self.rows, self.cols = int(layout), int(rows), int(cols)
        self.nbytes = int(self._lib.bitnet_b200_weight_nbytes(h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.bitnet_b200_weight_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
