"""Decode harness over Part 4 of the C ABI: a 2B-4T-shaped decoder step on device.

`ModelShape` + `synth_layer`/`synth_embedding`/`rope_tables` generate the random-init model
BASELINE.json's configs describe (numpy only, splitmix64 so any harness can regenerate the
same bytes). `Model` uploads it and steps it; nothing here computes on the CPU.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import ffi
from .device import LAYOUT_BITNET32_F16, LAYOUT_QK256


@dataclass
class ModelShape:
    hidden: int = 2560
    ffn: int = 6912
    n_layers: int = 30
    n_heads: int = 20
    n_kv_heads: int = 5
    head_dim: int = 128
    vocab: int = 128256
    max_ctx: int = 256
    batch: int = 1
    rms_eps: float = 1e-5
    layout: int = LAYOUT_QK256
    rope_theta: float = 500000.0

    @staticmethod
    def bitnet_2b4t(**kw):
        """BitNet b1.58-2B-4T (crates/bitnet-models/src/qk256_utils.rs:96-110)."""
        return ModelShape(**kw)

    @staticmethod
    def llama_8b_class(**kw):
        """8B-class ternary model of BASELINE.json configs[4] (hidden 4096, FFN 14336, 32 layers)."""
        d = dict(hidden=4096, ffn=14336, n_layers=32, n_heads=32, n_kv_heads=8, head_dim=128)
        d.update(kw)
        return ModelShape(**d)

    def row_bytes(self, cols):
        return -(-cols // 256) * 64 if self.layout == LAYOUT_QK256 else cols // 32 * 10

    def proj_shapes(self):
        qd, kd = self.n_heads * self.head_dim, self.n_kv_heads * self.head_dim
        return {"wq": (qd, self.hidden), "wk": (kd, self.hidden), "wv": (kd, self.hidden),
                "wo": (self.hidden, qd), "wgate": (self.ffn, self.hidden),
                "wup": (self.ffn, self.hidden), "wdown": (self.hidden, self.ffn)}

    def ternary_bytes(self):
        """Packed ternary weight bytes of one decode step (the algorithmic weight traffic)."""
        return self.n_layers * sum(r * self.row_bytes(c) for r, c in self.proj_shapes().values())


def _splitmix(seed, n):
    """n uint64 values of the counter-based splitmix64 stream `seed`."""
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + (np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def synth_bytes(n, seed):
    return _splitmix(seed, -(-n // 8)).view(np.uint8)[:n].copy()


def synth_uniform(n, seed, lo, hi):
    u = (_splitmix(seed, n) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    return (lo + (hi - lo) * u).astype(np.float32)


def synth_weight(shape, rows, cols, seed):
    """Random-init packed tensor: uniform code bytes; BitNet32-F16 scale = f16(u * K^-1/2)."""
    if shape.layout == LAYOUT_QK256:
        return synth_bytes(rows * shape.row_bytes(cols), seed)
    nb = rows * (cols // 32)
    codes = synth_bytes(nb * 8, seed).reshape(nb, 8)
    sc = (synth_uniform(nb, seed ^ 0x5CA1E, 0.5, 1.5) * np.float32(cols) ** np.float32(-0.5))
    sc = sc.astype(np.float16).view(np.uint8).reshape(nb, 2)
    return np.concatenate([codes, sc], axis=1).reshape(-1).copy()


def synth_layer(shape, layer, seed=0xB2000000):
    out = {}
    for i, (name, (r, c)) in enumerate(shape.proj_shapes().items()):
        out[name] = synth_weight(shape, r, c, seed + layer * 16 + i)
    out["attn_norm"] = synth_uniform(shape.hidden, seed + layer * 16 + 8, 0.5, 1.5)
    out["ffn_norm"] = synth_uniform(shape.hidden, seed + layer * 16 + 9, 0.5, 1.5)
    return out


def synth_embedding(shape, seed=0xE3B00000):
    """Tied embedding as f16 bits: U(-1, 1) rounded to f16 (exactly representable on both sides).

    Only the first min(vocab, 4096) rows are drawn from the generator; block j of 4096 rows is
    block 0 rotated by j columns (distinct rows, 30x cheaper to generate than 328 M draws)."""
    base_rows = min(shape.vocab, 4096)
    base = synth_uniform(base_rows * shape.hidden, seed, -1.0, 1.0).astype(np.float16)
    base = base.view(np.uint16).reshape(base_rows, shape.hidden)
    emb = np.empty((shape.vocab, shape.hidden), np.uint16)
    for j in range(-(-shape.vocab // base_rows)):
        lo, hi = j * base_rows, min(shape.vocab, (j + 1) * base_rows)
        emb[lo:hi] = np.roll(base[:hi - lo], j, axis=1)
    return emb.reshape(-1), synth_uniform(shape.hidden, seed + 1, 0.5, 1.5)


def rope_tables(shape):
    """cos/sin [max_ctx, head_dim/2] f32, computed in f64; both harnesses upload the SAME table."""
    half = shape.head_dim // 2
    inv_freq = shape.rope_theta ** (-np.arange(half, dtype=np.float64) * 2.0 / shape.head_dim)
    ang = np.arange(shape.max_ctx, dtype=np.float64)[:, None] * inv_freq[None, :]
    return np.cos(ang).astype(np.float32), np.sin(ang).astype(np.float32)


class Model:
    """bitnet_b200_model: upload once, then one C call per decode step."""

    def __init__(self, ctx, shape):
        self._lib = ffi.lib()
        self.ctx, self.shape = ctx, shape
        cfg = ffi.ModelConfig(shape.hidden, shape.ffn, shape.n_layers, shape.n_heads, shape.n_kv_heads,
                              shape.head_dim, shape.vocab, shape.max_ctx, shape.batch, shape.rms_eps,
                              shape.layout)
        h = C.c_void_p()
        ffi.check(self._lib.bitnet_b200_model_create(ctx._h, C.byref(cfg), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._lib.bitnet_b200_model_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_layer(self, layer, w):
        arrs = [np.ascontiguousarray(w[k], dtype=np.uint8) for k in
                ("wq", "wk", "wv", "wo", "wgate", "wup", "wdown")]
        norms = [np.ascontiguousarray(w[k], dtype=np.float32) for k in ("attn_norm", "ffn_norm")]
        for a, (name, (r, c)) in zip(arrs, self.shape.proj_shapes().items()):
            if a.size != r * self.shape.row_bytes(c):
                raise ValueError("%s: %d bytes, expected %d" % (name, a.size, r * self.shape.row_bytes(c)))
        ffi.check(self._lib.bitnet_b200_model_set_layer(self._h, layer, *[a.ctypes.data for a in arrs + norms]))

    def set_embedding(self, emb_f16, final_norm):
        emb = np.ascontiguousarray(emb_f16, dtype=np.uint16)
        fn = np.ascontiguousarray(final_norm, dtype=np.float32)
        if emb.size != self.shape.vocab * self.shape.hidden:
            raise ValueError("embedding size mismatch")
        ffi.check(self._lib.bitnet_b200_model_set_embedding(self._h, emb.ctypes.data, fn.ctypes.data))

    def set_rope(self, cos_t, sin_t):
        c = np.ascontiguousarray(cos_t, dtype=np.float32)
        s = np.ascontiguousarray(sin_t, dtype=np.float32)
        ffi.check(self._lib.bitnet_b200_model_set_rope(self._h, c.ctypes.data, s.ctypes.data))

    def load_synthetic(self, seed=0xB2000000):
        for l in range(self.shape.n_layers):
            self.set_layer(l, synth_layer(self.shape, l, seed))
        self.set_embedding(*synth_embedding(self.shape))
        self.set_rope(*rope_tables(self.shape))

    @classmethod
    def load_gguf(cls, ctx, path, max_ctx=256, batch=1):
        """Whole-model load from a GGUF file: tensor-name remap (`blk.N.attn_q.weight` -> layer slot, the reference's
        weight_mapper.rs:254-303,375-490), I2_S flavour detection per tensor, packed bytes uploaded as they lie on
        disk (loader.rs:1405-1521; no f32 dequantisation), norms / tied f16 embedding / RoPE tables from the
        metadata. Returns a ready Model; its `.shape` is read from the file."""
        from . import gguf
        r = gguf.GgufReader(path)
        try:
            fields, layers, emb, final_norm = gguf.load_model_tensors(r)
        finally:
            r.close()
        shape = ModelShape(max_ctx=max_ctx, batch=batch, **fields)
        m = cls(ctx, shape)
        for l, w in enumerate(layers):
            m.set_layer(l, w)
        m.set_embedding(emb, final_norm)
        m.set_rope(*rope_tables(shape))
        return m

    def reset(self):
        ffi.check(self._lib.bitnet_b200_model_reset(self._h))

    def decode(self, tokens, want_logits=False):
        """One greedy step through host buffers; returns next tokens (and logits)."""
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        nxt = np.zeros(self.shape.batch, dtype=np.int32)
        logits = np.zeros((self.shape.batch, self.shape.vocab), np.float32) if want_logits else None
        ffi.check(self._lib.bitnet_b200_model_decode(self._h, tok.ctypes.data, nxt.ctypes.data,
                                                     None if logits is None else logits.ctypes.data))
        return (nxt, logits) if want_logits else nxt

    def decode_device(self, n_steps):
        ffi.check(self._lib.bitnet_b200_model_decode_device(self._h, int(n_steps)))

    def set_tokens(self, tokens):
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        ffi.check(self._lib.bitnet_b200_model_set_tokens(self._h, tok.ctypes.data))

    def get_tokens(self):
        nxt = np.zeros(self.shape.batch, dtype=np.int32)
        ffi.check(self._lib.bitnet_b200_model_get_tokens(self._h, nxt.ctypes.data))
        return nxt

    def time_linears(self, reps=20):
        """(ms per pass, launches per pass) of only the ternary linears of one step."""
        ms, n = C.c_float(0), C.c_int(0)
        ffi.check(self._lib.bitnet_b200_model_time_linears(self._h, int(reps), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    @property
    def launches_per_step(self):
        return int(self._lib.bitnet_b200_model_launches_per_step(self._h))

    @property
    def weight_bytes(self):
        return int(self._lib.bitnet_b200_model_weight_bytes(self._h))
