"""ctypes binding of oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module (as the checker / the timed CPU baseline). The product
package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_LIB_PATH = os.path.join(HERE, "_ref", "libggml_quants_ref.so")

MAP_A = np.array([-2, -1, 1, 2], dtype=np.int8)   # i2s_qk256.rs:139-146
MAP_B = np.array([-2, -1, 0, 1], dtype=np.int8)   # quantization/src/utils.rs:76-91
MAP_C = np.array([0, 1, 0, -1], dtype=np.int8)    # cpu/quantized_matmul.rs:19-27


def build(force=False):
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    srcs = [os.path.join(HERE, f) for f in ("oracle.cpp", "model.cpp", "oracle.h")]
    stale = (not os.path.exists(LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale or not os.path.exists(REF_LIB_PATH):
        subprocess.run(["make", "-C", HERE], check=True, capture_output=True)


class OracleError(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_last_error.restype = C.c_char_p
        _lib.ref_code_to_f32.restype = C.c_float
        _lib.ref_code_to_f32.argtypes = [C.c_uint8]
        _lib.ref_gemv_qk256_row.restype = C.c_float
        _lib.ref_f16_to_f32.restype = C.c_float
        _lib.ref_f16_to_f32.argtypes = [C.c_uint16]
        _lib.ref_f32_to_f16.restype = C.c_uint16
        _lib.ref_f32_to_f16.argtypes = [C.c_float]
        for n in ("ref_i2s_expected_bytes", "ref_i2s_infer_block_size", "ref_qk256_tolerance_bytes"):
            getattr(_lib, n).restype = C.c_size_t
        _lib.ref_pack_i2s.restype = C.c_uint8
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _sz(v):
    return C.c_size_t(int(v))


def _chk(rc):
    if rc != 0:
        msg = lib().oracle_last_error()
        raise OracleError(msg.decode() if msg else "oracle error")


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


# ---- KernelProvider::matmul_i2s ---------------------------------------------------
def matmul_i2s(a, b, m, n, k, variant="fallback", c_len=None):
    a = _c(a, np.int8).ravel()
    b = _c(b, np.uint8).ravel()
    c = np.full(m * n if c_len is None else c_len, np.nan, dtype=np.float32)
    fn = getattr(lib(), "ref_matmul_i2s_" + variant)
    _chk(fn(_p(a), _sz(a.size), _p(b), _sz(b.size), _p(c), _sz(c.size), _sz(m), _sz(n), _sz(k)))
    return c.reshape(m, n)


def quantize(x, qtype=0, out=None, scales_len=None, out_len=None):
    x = _c(x, np.float32).ravel()
    block = {0: 32, 1: 64, 2: 128}.get(qtype, 32)
    if out is None:
        out = np.zeros((x.size + 3) // 4 if out_len is None else out_len, dtype=np.uint8)
    scales = np.zeros(-(-x.size // block) if scales_len is None else scales_len, dtype=np.float32)
    _chk(lib().ref_quantize(_p(x), _sz(x.size), _p(out), _sz(out.size), _p(scales),
                            _sz(scales.size), C.c_int(qtype)))
    return out, scales


def quantize_symmetric_i8(x, bits=8):
    x = _c(x, np.float32).ravel()
    q = np.zeros(x.size, dtype=np.int8)
    s = C.c_float(0)
    _chk(lib().ref_quantize_symmetric_i8(_p(x), _sz(x.size), C.c_int(bits), _p(q), C.byref(s)))
    return q, np.float32(s.value)


def quantize_rows_i8(x):
    """Per-token (row) absmax int8: quantize_symmetric_i8 on each row."""
    x = _c(x, np.float32)
    q = np.zeros(x.shape, dtype=np.int8)
    s = np.zeros(x.shape[0], dtype=np.float32)
    for i in range(x.shape[0]):
        q[i], s[i] = quantize_symmetric_i8(x[i])
    return q, s


# ---- QK256 ------------------------------------------------------------------------
def unpack_qk256_block(qs64):
    qs64 = _c(qs64, np.uint8)
    assert qs64.size == 64
    out = np.zeros(256, dtype=np.uint8)
    lib().ref_unpack_qk256_block(_p(qs64), _p(out))
    return out


def gemv_qk256_row(qs_row, x, cols):
    qs_row = _c(qs_row, np.uint8)
    x = _c(x, np.float32)
    return np.float32(lib().ref_gemv_qk256_row(_p(qs_row), _p(x), _sz(cols)))


def gemv_qk256(qs, x, rows, cols, stride, variant="dispatch", y_len=None, threads=0):
    qs = _c(qs, np.uint8).ravel()
    x = _c(x, np.float32).ravel()
    y = np.full(rows if y_len is None else y_len, np.nan, dtype=np.float32)
    name = {"dispatch": "ref_gemv_qk256", "scalar": "ref_gemv_qk256_scalar",
            "avx2": "ref_gemv_qk256_avx2", "mt": "ref_gemv_qk256_mt"}[variant]
    args = [_p(qs), _sz(qs.size), _p(x), _sz(x.size), _p(y), _sz(y.size), _sz(rows), _sz(cols),
            _sz(stride)]
    if variant == "mt":
        args.append(C.c_int(threads))
    _chk(getattr(lib(), name)(*args))
    return y


def qk256_new_check(rows, cols, qs_len):
    stride = C.c_size_t(0)
    _chk(lib().ref_i2s_qk256_new_check(_sz(rows), _sz(cols), _sz(qs_len), C.byref(stride)))
    return stride.value


def qk256_gemv_scalar_legacy(rows, cols, packed, scales, act):
    packed = _c(packed, np.uint8)
    scales = _c(scales, np.float32)
    act = _c(act, np.float32)
    out = np.zeros(rows, dtype=np.float32)
    _chk(lib().ref_qk256_gemv_scalar_legacy(_p(out), _sz(rows), _sz(cols), _p(packed),
                                            _sz(packed.size), _p(scales), _sz(scales.size), _p(act)))
    return out


# ---- BitNet32-F16 -------------------------------------------------------------------
def f16_to_f32(bits):
    return np.float32(lib().ref_f16_to_f32(C.c_uint16(int(bits))))


def f32_to_f16(f):
    return int(lib().ref_f32_to_f16(C.c_float(float(f))))


def i2s_dequant_block(qbits, n, scale_bits, inv=False, k=1.0):
    qbits = _c(qbits, np.uint8)
    dst = np.zeros(n, dtype=np.float32)
    lib().ref_i2s_dequant_block(_p(dst), _p(qbits), _sz(n), C.c_uint16(int(scale_bits)),
                                C.c_int(int(inv)), C.c_float(k))
    return dst


def i2s_expected_bytes(rows, cols, block):
    return lib().ref_i2s_expected_bytes(_sz(rows), _sz(cols), _sz(block))


def i2s_infer_block_size(nbytes, rows, cols):
    b = lib().ref_i2s_infer_block_size(_sz(nbytes), _sz(rows), _sz(cols))
    return None if b == 0 else b


def dequantize_i2s(data, rows, cols, block=None, inv=False, k=1.0):
    data = _c(data, np.uint8).ravel()
    out = np.zeros(rows * cols, dtype=np.float32)
    if block is None:
        _chk(lib().ref_dequantize_i2s(_p(data), _sz(data.size), _sz(rows), _sz(cols),
                                      C.c_int(int(inv)), C.c_float(k), _p(out)))
    else:
        _chk(lib().ref_dequantize_i2s_with_block(_p(data), _sz(data.size), _sz(rows), _sz(cols),
                                                 _sz(block), C.c_int(int(inv)), C.c_float(k),
                                                 _p(out)))
    return out.reshape(rows, cols)


# ---- layout detection -------------------------------------------------------------
FLAVORS = {0: "BitNet32F16", 1: "Split32WithSibling", 2: "GgmlQk256NoScale", -1: None}


def qk256_tolerance_bytes(expected):
    return lib().ref_qk256_tolerance_bytes(_sz(expected))


def detect_i2s_flavor(available, has_sibling, nelems, strict=False):
    return FLAVORS[lib().ref_detect_i2s_flavor(_sz(available), C.c_int(int(has_sibling)),
                                               _sz(nelems), C.c_int(int(strict)))]


# ---- packed + scaled matmul (map C) -------------------------------------------------
def pack_i2s(vals):
    v = _c(vals, np.int8)
    assert v.size == 4
    return int(lib().ref_pack_i2s(_p(v)))


def i2s_matmul(act, w, scales, m, n, k, block_size, variant="f32", out_len=None):
    act = _c(act, np.float32).ravel()
    w = _c(w, np.uint8).ravel()
    scales = _c(scales, np.float32).ravel()
    out = np.full(m * n if out_len is None else out_len, np.nan, dtype=np.float32)
    name = {"f32": "ref_i2s_matmul_f32", "blocked": "ref_i2s_matmul_blocked",
            "dequant": "ref_dequantize_and_matmul"}[variant]
    _chk(getattr(lib(), name)(_p(act), _sz(act.size), _p(w), _sz(w.size), _p(scales),
                              _sz(scales.size), _p(out), _sz(out.size), _sz(m), _sz(n), _sz(k),
                              _sz(block_size)))
    return out[:m * n].reshape(m, n)


def unpack_2bit_values(packed, output_len):
    packed = _c(packed, np.uint8)
    out = np.zeros(output_len, dtype=np.int8)
    lib().ref_unpack_2bit_values(_p(packed), _sz(packed.size), _sz(output_len), _p(out))
    return out


def pack_2bit_values(vals):
    vals = _c(vals, np.int8)
    out = np.zeros((vals.size + 3) // 4, dtype=np.uint8)
    lib().ref_pack_2bit_values(_p(vals), _sz(vals.size), _p(out))
    return out


# ---- RMSNorm ----------------------------------------------------------------------
def rms_norm(x, gamma, eps=1e-5):
    x = _c(x, np.float32)
    gamma = _c(gamma, np.float32)
    out = np.zeros(x.size, dtype=np.float32)
    _chk(lib().ref_rms_norm(_p(x), _sz(x.size), _p(gamma), _sz(gamma.size), C.c_float(eps), _p(out)))
    return out.reshape(x.shape)


def fused_rmsnorm_linear(x, weight, gamma, eps):
    x = _c(x, np.float32)
    weight = _c(weight, np.float32)
    gamma = _c(gamma, np.float32)
    out = np.zeros(weight.size // x.size, dtype=np.float32)
    _chk(lib().ref_fused_rmsnorm_linear(_p(x), _sz(x.size), _p(weight), _sz(weight.size),
                                        _p(gamma), C.c_float(eps), _p(out)))
    return out


# ---- composed int8 linears ----------------------------------------------------------
def linear_qk256_i8(qs, rows, cols, q, act_scale, lut=MAP_A):
    qs = _c(qs, np.uint8).ravel()
    q = _c(q, np.int8)
    m = q.shape[0]
    act_scale = _c(act_scale, np.float32)
    lut = _c(lut, np.int8)
    stride = -(-cols // 256) * 64
    acc = np.zeros((m, rows), dtype=np.int32)
    y = np.zeros((m, rows), dtype=np.float32)
    _chk(lib().oracle_linear_qk256_i8(_p(qs), _sz(rows), _sz(cols), _sz(stride), _p(q),
                                      _p(act_scale), _sz(m), _p(lut), _p(acc), _p(y)))
    return acc, y


def linear_bitnet32_i8(data, rows, cols, q, act_scale, lut=MAP_A, scale_policy=0, want_blocks=False):
    data = _c(data, np.uint8).ravel()
    q = _c(q, np.int8)
    m = q.shape[0]
    act_scale = _c(act_scale, np.float32)
    lut = _c(lut, np.int8)
    y = np.zeros((m, rows), dtype=np.float32)
    acc = np.zeros((m, rows, cols // 32), dtype=np.int32) if want_blocks else None
    _chk(lib().oracle_linear_bitnet32_i8(_p(data), _sz(rows), _sz(cols), _p(q), _p(act_scale),
                                         _sz(m), _p(lut), C.c_int(scale_policy),
                                         _p(acc) if want_blocks else None, _p(y)))
    return (acc, y) if want_blocks else y


def linear_bitnet32_dequant(data, rows, cols, q, act_scale):
    data = _c(data, np.uint8).ravel()
    q = _c(q, np.int8)
    m = q.shape[0]
    act_scale = _c(act_scale, np.float32)
    y = np.zeros((m, rows), dtype=np.float32)
    _chk(lib().oracle_linear_bitnet32_dequant(_p(data), _sz(rows), _sz(cols), _p(q),
                                              _p(act_scale), _sz(m), _p(y)))
    return y


# ---- synthetic data (splitmix64) ------------------------------------------
def fill_bytes(n, seed):
    out = np.zeros(n, dtype=np.uint8)
    lib().oracle_fill_bytes(_p(out), _sz(n), C.c_uint64(seed))
    return out


def fill_normal(n, seed, sigma=1.0):
    out = np.zeros(n, dtype=np.float32)
    lib().oracle_fill_normal(_p(out), _sz(n), C.c_uint64(seed), C.c_float(sigma))
    return out


def fill_uniform(n, seed, lo, hi):
    out = np.zeros(n, dtype=np.float32)
    lib().oracle_fill_uniform(_p(out), _sz(n), C.c_uint64(seed), C.c_float(lo), C.c_float(hi))
    return out


def make_bitnet32(rows, cols, seed):
    """Random BitNet32-F16 tensor: splitmix code bytes, f16 scale = f16(u * K^-1/2), u~U[0.5,1.5)."""
    nb = rows * (cols // 32)
    codes = fill_bytes(nb * 8, seed).reshape(nb, 8)
    u = fill_uniform(nb, seed ^ 0x5CA1E, 0.5, 1.5) * np.float32(cols) ** np.float32(-0.5)
    sc = u.astype(np.float16).view(np.uint8).reshape(nb, 2)
    return np.concatenate([codes, sc], axis=1).reshape(-1).copy()


# ---- reference-built C (oracle/_ref) --------------------------------------------------
def ref_lib():
    if not os.path.exists(REF_LIB_PATH):
        return None
    l = C.CDLL(REF_LIB_PATH)
    l.bitnet_dequantize_row_iq2_s.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    l.bitnet_dequantize_row_iq2_s.restype = None
    return l


# ---- 2B-4T-shaped decoder (oracle/model.cpp) ------------------------------------------
class _ModelConfig(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("hidden", "ffn", "n_layers", "n_heads", "n_kv_heads", "head_dim",
                                        "vocab", "max_ctx")] + [("rms_eps", C.c_float), ("layout", C.c_int)]


class Model:
    """CPU oracle model; `shape` is any object with the ModelShape fields."""

    def __init__(self, shape, n_seqs=1, threads=None):
        self._lib = lib()
        self._lib.oracle_model_create.restype = C.c_void_p
        self._lib.oracle_model_create.argtypes = [C.POINTER(_ModelConfig), C.c_int]
        self._lib.oracle_model_destroy.argtypes = [C.c_void_p]
        self._lib.oracle_model_set_layer.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 9
        self._lib.oracle_model_set_embedding.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self._lib.oracle_model_set_rope.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self._lib.oracle_model_reset.argtypes = [C.c_void_p]
        self._lib.oracle_model_step.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        self.shape = shape
        self.threads = threads or os.cpu_count() or 1
        cfg = _ModelConfig(shape.hidden, shape.ffn, shape.n_layers, shape.n_heads, shape.n_kv_heads,
                           shape.head_dim, shape.vocab, shape.max_ctx, shape.rms_eps, shape.layout)
        self._h = self._lib.oracle_model_create(C.byref(cfg), n_seqs)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.oracle_model_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_layer(self, layer, w):
        arrs = [_c(w[k], np.uint8) for k in ("wq", "wk", "wv", "wo", "wgate", "wup", "wdown")]
        arrs += [_c(w[k], np.float32) for k in ("attn_norm", "ffn_norm")]
        _chk(self._lib.oracle_model_set_layer(self._h, layer, *[_p(a) for a in arrs]))

    def set_embedding(self, emb_f16, final_norm):
        e, f = _c(emb_f16, np.uint16), _c(final_norm, np.float32)
        _chk(self._lib.oracle_model_set_embedding(self._h, _p(e), _p(f)))

    def set_rope(self, cos_t, sin_t):
        c, s = _c(cos_t, np.float32), _c(sin_t, np.float32)
        _chk(self._lib.oracle_model_set_rope(self._h, _p(c), _p(s)))

    def reset(self):
        self._lib.oracle_model_reset(self._h)

    def set_threads(self, n):
        """host threads the rows of every projection are split over (1 = the reference's serial loops)"""
        self.threads = max(1, int(n))

    def step(self, token, seq=0, want_logits=False):
        logits = np.zeros(self.shape.vocab, np.float32) if want_logits else None
        t = self._lib.oracle_model_step(self._h, seq, int(token), None if logits is None else _p(logits),
                                        None, self.threads)
        if t < 0:
            raise OracleError("oracle_model_step failed")
        return (t, logits) if want_logits else t
