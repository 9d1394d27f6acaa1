"""ctypes loader for libbitnet_b200.so (C ABI: include/bitnet_b200.h)."""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
# BITNET_B200_LIB: load another build of the same library (diagnostics: the `make trace` build)
LIB_PATH = os.environ.get("BITNET_B200_LIB") or os.path.join(HERE, "libbitnet_b200.so")
CSRC = os.path.join(HERE, "csrc")

# every symbol include/bitnet_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "bitnet_cpp_init", "bitnet_cpp_cleanup", "bitnet_cpp_is_available", "bitnet_cpp_matmul_i2s",
    "bitnet_cpp_quantize", "bitnet_cpp_get_last_error", "bitnet_b200_gemv_qk256",
    "bitnet_b200_qk256_gemv", "bitnet_b200_qk256_gemv_scalar",
    "bitnet_b200_ctx_create", "bitnet_b200_ctx_set_stream", "bitnet_b200_ctx_synchronize",
    "bitnet_b200_ctx_destroy", "bitnet_b200_weight_create", "bitnet_b200_weight_destroy",
    "bitnet_b200_weight_rows", "bitnet_b200_weight_cols", "bitnet_b200_weight_nbytes",
    "bitnet_b200_actquant_i8", "bitnet_b200_rmsnorm_actquant", "bitnet_b200_linear_i8", "bitnet_b200_linear_i8_swiglu",
    "bitnet_b200_linear_i8_static_scratch",
    "bitnet_b200_linear_f32", "bitnet_b200_fused_linear", "bitnet_b200_i2s_matmul_f32", "bitnet_b200_version",
    "bitnet_b200_launch_count",
    "bitnet_b200_last_linear_kernel",
    "bitnet_b200_model_create", "bitnet_b200_model_destroy", "bitnet_b200_model_set_layer",
    "bitnet_b200_model_set_embedding", "bitnet_b200_model_set_rope", "bitnet_b200_model_reset",
    "bitnet_b200_model_decode", "bitnet_b200_model_decode_device", "bitnet_b200_model_set_tokens",
    "bitnet_b200_model_get_tokens", "bitnet_b200_model_launches_per_step",
    "bitnet_b200_model_time_linears",
    "bitnet_b200_model_weight_bytes",
    "bitnet_b200_row_absmax", "bitnet_b200_actquant_i8_absmax", "bitnet_b200_residual_acc",
    "bitnet_b200_rope_attn_decode",
    "bitnet_b200_tp_xchg_sum_rmsnorm_actquant", "bitnet_b200_tp_xchg_max_actquant",
]


class TpXchg(C.Structure):
    """bitnet_b200_tp_xchg: one cross-GPU exchange point (include/bitnet_b200.h, Part 5)"""
    _fields_ = [("peer_data", C.c_void_p * 8), ("peer_flags", C.c_void_p * 8), ("counters", C.c_void_p),
                ("n_ranks", C.c_int), ("rank", C.c_int), ("slot", C.c_int), ("phase", C.c_int)]


class ModelConfig(C.Structure):
    """bitnet_b200_model_config"""
    _fields_ = [(n, C.c_int) for n in ("hidden", "ffn", "n_layers", "n_heads", "n_kv_heads", "head_dim",
                                        "vocab", "max_ctx", "batch")] + [("rms_eps", C.c_float),
                                                                         ("layout", C.c_int)]


class B200Error(RuntimeError):
    """An error reported by libbitnet_b200.so (message = bitnet_cpp_get_last_error())."""


def build(force=False, extra=""):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC]
    if force:
        cmd.append("-B")
    if extra:
        cmd.append("EXTRA=" + extra)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise B200Error("building libbitnet_b200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


_lib = None


def lib():
    """The loaded library. Fails loudly when it has not been built -- no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(
                "libbitnet_b200.so is missing (%s): build it with __graft_entry__.build() or "
                "`make -C bitnet-rs_b200/csrc`; there is no CPU fallback" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        vp, sz, i32 = C.c_void_p, C.c_size_t, C.c_int
        l.bitnet_cpp_init.restype = i32
        l.bitnet_cpp_cleanup.restype = None
        l.bitnet_cpp_is_available.restype = i32
        l.bitnet_cpp_get_last_error.restype = C.c_char_p
        l.bitnet_cpp_matmul_i2s.argtypes = [vp, vp, vp, i32, i32, i32]
        l.bitnet_cpp_quantize.argtypes = [vp, i32, vp, i32, vp, i32, i32]
        l.bitnet_b200_gemv_qk256.argtypes = [vp, sz, vp, sz, vp, sz, sz, sz, sz]
        l.bitnet_b200_ctx_create.argtypes = [i32, vp, C.POINTER(vp)]
        l.bitnet_b200_ctx_set_stream.argtypes = [vp, vp]
        l.bitnet_b200_ctx_synchronize.argtypes = [vp]
        l.bitnet_b200_ctx_destroy.argtypes = [vp]
        l.bitnet_b200_ctx_destroy.restype = None
        l.bitnet_b200_weight_create.argtypes = [vp, vp, sz, i32, sz, sz, vp, i32, C.POINTER(vp)]
        l.bitnet_b200_weight_destroy.argtypes = [vp]
        l.bitnet_b200_weight_destroy.restype = None
        for n in ("rows", "cols", "nbytes"):
            f = getattr(l, "bitnet_b200_weight_" + n)
            f.argtypes = [vp]
            f.restype = sz
        l.bitnet_b200_actquant_i8.argtypes = [vp, vp, i32, i32, vp, vp]
        l.bitnet_b200_rmsnorm_actquant.argtypes = [vp, vp, vp, C.c_float, i32, i32, vp, vp, vp]
        l.bitnet_b200_linear_i8.argtypes = [vp, vp, vp, vp, i32, vp, vp]
        l.bitnet_b200_linear_i8_swiglu.argtypes = [vp, vp, vp, vp, i32, vp]
        l.bitnet_b200_linear_i8_static_scratch.argtypes = [vp, vp, vp, vp, i32, vp]
        l.bitnet_b200_linear_f32.argtypes = [vp, vp, vp, i32, vp]
        l.bitnet_b200_fused_linear.argtypes = [vp, vp, vp, vp, C.c_float, i32, i32, vp, vp, vp, vp]
        l.bitnet_b200_i2s_matmul_f32.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32]
        l.bitnet_b200_model_create.argtypes = [vp, C.POINTER(ModelConfig), C.POINTER(vp)]
        l.bitnet_b200_model_destroy.argtypes = [vp]
        l.bitnet_b200_model_destroy.restype = None
        l.bitnet_b200_model_set_layer.argtypes = [vp, i32] + [vp] * 9
        l.bitnet_b200_model_set_embedding.argtypes = [vp, vp, vp]
        l.bitnet_b200_model_set_rope.argtypes = [vp, vp, vp]
        l.bitnet_b200_model_reset.argtypes = [vp]
        l.bitnet_b200_model_decode.argtypes = [vp, vp, vp, vp]
        l.bitnet_b200_model_decode_device.argtypes = [vp, i32]
        l.bitnet_b200_model_set_tokens.argtypes = [vp, vp]
        l.bitnet_b200_model_get_tokens.argtypes = [vp, vp]
        l.bitnet_b200_model_time_linears.argtypes = [vp, i32, C.POINTER(C.c_float), C.POINTER(i32)]
        l.bitnet_b200_model_launches_per_step.argtypes = [vp]
        l.bitnet_b200_model_launches_per_step.restype = C.c_uint64
        l.bitnet_b200_model_weight_bytes.argtypes = [vp]
        l.bitnet_b200_model_weight_bytes.restype = C.c_uint64
        l.bitnet_b200_row_absmax.argtypes = [vp, vp, i32, i32, i32, vp, vp]
        l.bitnet_b200_actquant_i8_absmax.argtypes = [vp, vp, vp, i32, i32, vp, vp]
        l.bitnet_b200_residual_acc.argtypes = [vp, vp, vp, vp, i32, i32]
        l.bitnet_b200_rope_attn_decode.argtypes = [vp] * 7 + [i32] * 5 + [vp]
        l.bitnet_b200_tp_xchg_sum_rmsnorm_actquant.argtypes = [vp, C.POINTER(TpXchg), vp, vp, vp, C.c_float, i32, i32, vp, vp, vp]
        l.bitnet_b200_tp_xchg_max_actquant.argtypes = [vp, C.POINTER(TpXchg), vp, i32, i32, vp, vp]
        for fn in (l.bitnet_b200_qk256_gemv, l.bitnet_b200_qk256_gemv_scalar):
            fn.argtypes = [vp, sz, sz, sz, vp, sz, vp, sz, vp, sz]
        l.bitnet_b200_version.restype = C.c_char_p
        l.bitnet_b200_launch_count.restype = C.c_uint64
        l.bitnet_b200_last_linear_kernel.restype = C.c_char_p
        _lib = l
    return _lib


def last_error():
    m = lib().bitnet_cpp_get_last_error()
    return m.decode() if m else None


def check(rc):
    if rc != 0:
        raise B200Error(last_error() or "unknown libbitnet_b200 error")


def launch_count():
    return int(lib().bitnet_b200_launch_count())


def last_linear_kernel():
    """the kernel the calling thread's last linear_i8 ran (diagnostics)"""
    return lib().bitnet_b200_last_linear_kernel().decode()
