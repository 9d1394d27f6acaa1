"""GGUF I2_S tensor reader + layout (flavour) detection: straight from a `ggml-model-i2_s.gguf` on
disk to device-resident packed weights, in the on-disk byte layout.

Mirrors, on the host, what the reference does in
  crates/bitnet-models/src/formats/gguf/reader.rs        (header / metadata / tensor-info parsing,
                                                          data start = align_up(kv_end, general.alignment))
  crates/bitnet-models/src/formats/gguf/types.rs:764-845  (I2SFlavor, tensor type 36 = I2_S)
  crates/bitnet-models/src/formats/gguf/types.rs:868-1066 (detect_i2s_flavor: by available byte count)
  crates/bitnet-quantization/src/lib.rs:86-90             (qk256_tolerance_bytes)
  crates/bitnet-models/src/formats/gguf/loader.rs:1405-1521 (the I2_S branch of the loader)
but keeps the bytes packed: no dequantisation to f32 happens anywhere. numpy only; no compute.
"""
import enum
import mmap
import struct

import numpy as np

from .device import LAYOUT_BITNET32_F16, LAYOUT_QK256

GGUF_MAGIC = b"GGUF"
GGML_TYPE_F32, GGML_TYPE_F16, GGML_TYPE_I2_S = 0, 1, 36       # types.rs:689-705
DEFAULT_ALIGNMENT = 32


class I2SFlavor(enum.IntEnum):
    """types.rs:764-806"""
    BitNet32F16 = 0          # 32-element blocks, 8 B codes + 2 B f16 scale inline (10 B per block)
    Split32WithSibling = 1   # 32-element blocks, 8 B codes, scales in a sibling tensor
    GgmlQk256NoScale = 2     # 256-element blocks, 64 B codes, no scale


def qk256_tolerance_bytes(expected):
    """0.1 % of the expected size, at least 8 bytes (crates/bitnet-quantization/src/lib.rs:86-90)."""
    import math
    return max(8, int(math.ceil(expected * 0.001)))


def detect_i2s_flavor(available, has_scale_sibling, nelems, strict=False, name="?"):
    """types.rs:868-1066: exact byte-count matches first (QK256 > BitNet32F16 > Split32 + sibling), then
    matches within the tolerance (QK256 > Split32 + sibling > BitNet32F16 > Split32)."""
    blocks32, blocks256 = -(-nelems // 32), -(-nelems // 256)
    split_need, inline_need, qk256_need = blocks32 * 8, blocks32 * 10, blocks256 * 64
    tol = 8 if strict else qk256_tolerance_bytes(min(split_need, qk256_need))
    d_split, d_inline, d_qk = abs(available - split_need), abs(available - inline_need), abs(available - qk256_need)
    if d_qk == 0:
        return I2SFlavor.GgmlQk256NoScale
    if d_inline == 0:
        return I2SFlavor.BitNet32F16
    if d_split == 0 and has_scale_sibling:
        return I2SFlavor.Split32WithSibling
    if d_qk <= tol:
        return I2SFlavor.GgmlQk256NoScale
    if has_scale_sibling and d_split <= tol:
        return I2SFlavor.Split32WithSibling
    if d_inline <= tol:
        return I2SFlavor.BitNet32F16
    if d_split <= tol:
        return I2SFlavor.Split32WithSibling
    raise ValueError("I2_S '%s': no valid flavor detected (available=%d, split_need=%d, inline_need=%d, qk256_need=%d)"
                     % (name, available, split_need, inline_need, qk256_need))


class TensorInfo:
    def __init__(self, name, shape, ggml_type, offset):
        self.name, self.shape, self.ggml_type, self.offset = name, tuple(shape), ggml_type, offset
        self.size = 0   # bytes available up to the next tensor / end of file (reader.rs: info.size)

    @property
    def nelems(self):
        n = 1
        for d in self.shape:
            n *= d
        return n


_SCALAR = {0: "<B", 1: "<b", 2: "<H", 3: "<h", 4: "<I", 5: "<i", 6: "<f", 7: "<?", 10: "<Q", 11: "<q", 12: "<d"}


class GgufReader:
    """Header, metadata and tensor table of a GGUF v2/v3 file (memory-mapped; tensor bytes are views)."""

    def __init__(self, path):
        self.path = path
        self._f = open(path, "rb")
        self._mm = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)
        self._pos = 0
        if self._take(4) != GGUF_MAGIC:
            raise ValueError("not a GGUF file (bad magic)")
        self.version = self._u32()
        if self.version not in (2, 3):
            raise ValueError("unsupported GGUF version %d" % self.version)
        n_tensors, n_kv = self._u64(), self._u64()
        self.metadata = {}
        for _ in range(n_kv):
            key = self._str()
            self.metadata[key] = self._value(self._u32())
        self.tensors = {}
        order = []
        for _ in range(n_tensors):
            name = self._str()
            nd = self._u32()
            dims = [self._u64() for _ in range(nd)]
            t = self._u32()
            off = self._u64()
            info = TensorInfo(name, dims[::-1], t, off)     # GGUF lists the innermost dimension first
            self.tensors[name] = info
            order.append(info)
        self.alignment = int(self.metadata.get("general.alignment", DEFAULT_ALIGNMENT)) or DEFAULT_ALIGNMENT
        self.data_start = -(-self._pos // self.alignment) * self.alignment
        end = len(self._mm) - self.data_start
        by_off = sorted(order, key=lambda i: i.offset)
        for i, info in enumerate(by_off):
            nxt = by_off[i + 1].offset if i + 1 < len(by_off) else end
            info.size = nxt - info.offset
            if info.offset > end or info.size < 0:
                raise ValueError("tensor '%s' data extends beyond file bounds" % info.name)

    # ---- primitive readers ------------------------------------------------------------------
    def _take(self, n):
        b = self._mm[self._pos:self._pos + n]
        if len(b) != n:
            raise ValueError("truncated GGUF file")
        self._pos += n
        return b

    def _u32(self):
        return struct.unpack("<I", self._take(4))[0]

    def _u64(self):
        return struct.unpack("<Q", self._take(8))[0]

    def _str(self):
        return self._take(self._u64()).decode("utf-8", errors="replace")

    def _value(self, t):
        if t in _SCALAR:
            fmt = _SCALAR[t]
            return struct.unpack(fmt, self._take(struct.calcsize(fmt)))[0]
        if t == 8:
            return self._str()
        if t == 9:
            et, n = self._u32(), self._u64()
            return [self._value(et) for _ in range(n)]
        raise ValueError("unknown GGUF metadata value type %d" % t)

    # ---- tensors ----------------------------------------------------------------------------
    def tensor_bytes(self, name, nbytes=None):
        info = self.tensors[name]
        n = info.size if nbytes is None else nbytes
        start = self.data_start + info.offset
        return np.frombuffer(self._mm, dtype=np.uint8, count=n, offset=start)

    def i2s_tensors(self):
        return [i for i in self.tensors.values() if i.ggml_type == GGML_TYPE_I2_S]

    def i2s_flavor(self, name, strict=False):
        info = self.tensors[name]
        if info.ggml_type != GGML_TYPE_I2_S:
            raise ValueError("tensor '%s' is not I2_S (type %d)" % (name, info.ggml_type))
        sibling = any(n in self.tensors for n in (name + ".scale", name + "_scale", name.replace(".weight", ".scale")) if n != name)
        return detect_i2s_flavor(info.size, sibling, info.nelems, strict, name)

    def upload_i2s(self, ctx, name, strict=False):
        """The packed bytes of an I2_S tensor, uploaded as they lie on disk -> device.Weight.
        shape = [rows = out_features, cols = in_features] (crates/bitnet-models/src/qk256_utils.rs:19-56)."""
        info = self.tensors[name]
        if len(info.shape) != 2:
            raise ValueError("I2_S tensor '%s' is not a matrix: shape %r" % (name, info.shape))
        rows, cols = info.shape
        flavor = self.i2s_flavor(name, strict)
        if flavor == I2SFlavor.GgmlQk256NoScale:
            need = rows * (-(-cols // 256) * 64)          # row-padded; the file may hold nelems-based bytes (+-tolerance)
            layout = LAYOUT_QK256
        elif flavor == I2SFlavor.BitNet32F16:
            if cols % 32:
                raise ValueError("BitNet32-F16 tensor '%s': cols %d is not a multiple of 32" % (name, cols))
            need = rows * (cols // 32) * 10
            layout = LAYOUT_BITNET32_F16
        else:
            raise ValueError("I2_S '%s': the split layout with a sibling scale tensor is not produced by the "
                             "production loaders (loader.rs:1422) and is not uploaded" % name)
        if info.size < need - qk256_tolerance_bytes(need):
            raise ValueError("I2_S '%s': %d bytes on disk, %d needed for [%d, %d]" % (name, info.size, need, rows, cols))
        data = np.zeros(need, np.uint8)
        n = min(need, info.size)
        data[:n] = self.tensor_bytes(name, n)
        return ctx.weight(data, layout, rows, cols), flavor

    def close(self):
        self._mm.close()
        self._f.close()


# --------------------------------------------------------------------------------------------
# Tensor-name remap: GGUF / vendor names -> the transformer's canonical names
# (crates/bitnet-models/src/weight_mapper.rs:113-251 normalize_vendor_key, :375-490 map_tensor_name)
# --------------------------------------------------------------------------------------------
import re

_EMBED = {"token_embd.weight", "tok_embeddings.weight", "model.embed_tokens.weight", "transformer.wte.weight",
          "transformer.word_embeddings.weight", "embeddings.word_embeddings.weight", "embed.weight", "embedding.weight",
          "word_embeddings.weight"}
_LM_HEAD = {"output.weight", "lm_head.weight", "model.lm_head.weight", "generator.weight", "transformer.lm_head.weight",
            "language_model_head.weight", "head.weight", "cls.weight"}
_FINAL_NORM = {"output_norm.weight", "norm.weight", "model.norm.weight", "transformer.ln_f.weight", "ln_f.weight",
               "final_norm.weight", "final_layernorm.weight", "final_rmsnorm.weight"}
_BLK = {  # weight_mapper.rs:417-447
    "attn_q.weight": "attention.q_proj.weight", "attn_k.weight": "attention.k_proj.weight",
    "attn_v.weight": "attention.v_proj.weight", "attn_output.weight": "attention.o_proj.weight",
    "attn_o.weight": "attention.o_proj.weight", "attn_norm.weight": "attention_norm.weight",
    "attn_sub_norm.weight": "attention.sub_layernorm.weight",
    "ffn_gate.weight": "feed_forward.gate_proj.weight", "ffn_gate_inp.weight": "feed_forward.gate_proj.weight",
    "ffn_up.weight": "feed_forward.up_proj.weight", "ffn_up_proj.weight": "feed_forward.up_proj.weight",
    "ffn_down.weight": "feed_forward.down_proj.weight", "ffn_down_proj.weight": "feed_forward.down_proj.weight",
    "ffn_norm.weight": "post_attention_layernorm.weight", "ffn_sub_norm.weight": "feed_forward.sub_layernorm.weight",
}
_LAYERS = {  # weight_mapper.rs:452-490
    "attention.wq.weight": "attention.q_proj.weight", "self_attn.q_proj.weight": "attention.q_proj.weight",
    "attention.wk.weight": "attention.k_proj.weight", "self_attn.k_proj.weight": "attention.k_proj.weight",
    "attention.wv.weight": "attention.v_proj.weight", "self_attn.v_proj.weight": "attention.v_proj.weight",
    "attention.wo.weight": "attention.o_proj.weight", "self_attn.o_proj.weight": "attention.o_proj.weight",
    "attention.q_proj.weight": "attention.q_proj.weight", "attention.k_proj.weight": "attention.k_proj.weight",
    "attention.v_proj.weight": "attention.v_proj.weight", "attention.o_proj.weight": "attention.o_proj.weight",
    "attention_norm.weight": "attention_norm.weight", "input_layernorm.weight": "attention_norm.weight",
    "ffn_norm.weight": "post_attention_layernorm.weight", "post_attention_layernorm.weight": "post_attention_layernorm.weight",
    "feed_forward.w1.weight": "feed_forward.gate_proj.weight", "mlp.gate_proj.weight": "feed_forward.gate_proj.weight",
    "feed_forward.gate_proj.weight": "feed_forward.gate_proj.weight",
    "feed_forward.w3.weight": "feed_forward.up_proj.weight", "mlp.up_proj.weight": "feed_forward.up_proj.weight",
    "feed_forward.up_proj.weight": "feed_forward.up_proj.weight",
    "feed_forward.w2.weight": "feed_forward.down_proj.weight", "mlp.down_proj.weight": "feed_forward.down_proj.weight",
    "feed_forward.down_proj.weight": "feed_forward.down_proj.weight",
}


def map_tensor_name(name):
    """GGUF / vendor tensor name -> canonical `layers.N.attention.q_proj.weight` style name, or None when the
    tensor is not one the transformer uses (weight_mapper.rs:375-490)."""
    if name in _EMBED:
        return "embed_tokens.weight"
    if name in _LM_HEAD:
        return "lm_head.weight"
    if name in _FINAL_NORM:
        return "final_norm.weight"
    m = re.match(r"^blk\.(\d+)\.(.+)$", name)
    if m:
        comp = _BLK.get(m.group(2))
        return None if comp is None else "layers.%s.%s" % (m.group(1), comp)
    m = re.match(r"^(?:model\.)?layers\.(\d+)\.(.+)$", name)
    if m:
        comp = _LAYERS.get(m.group(2))
        return None if comp is None else "layers.%s.%s" % (m.group(1), comp)
    return None


def qk256_key(name):
    """Key under which the loader files the raw QK256 bytes of tensor `name` (loader.rs:1507)."""
    return name + ".qk256_qs"


def expected_qk256_shape(rows, cols):
    """[rows, ceil(cols/256) * 64] u8 (crates/bitnet-models/src/qk256_utils.rs:19-56)."""
    return rows, -(-cols // 256) * 64


# --------------------------------------------------------------------------------------------
# A minimal GGUF v3 writer: synthetic models for the loader tests and the bench (there are no real checkpoints
# in the build image). Same container rules as the reader above.
# --------------------------------------------------------------------------------------------
def write_gguf(path, metadata, tensors, alignment=DEFAULT_ALIGNMENT):
    """metadata: {key: int | float | str}; tensors: [(name, shape (outermost first), ggml_type, bytes-like)]."""
    def s(x):
        b = x.encode("utf-8")
        return struct.pack("<Q", len(b)) + b

    out = bytearray(GGUF_MAGIC + struct.pack("<IQQ", 3, len(tensors), len(metadata) + 1))
    meta = dict(metadata)
    meta["general.alignment"] = int(alignment)
    for k, v in meta.items():
        out += s(k)
        if isinstance(v, bool):
            out += struct.pack("<I?", 7, v)
        elif isinstance(v, int):
            out += struct.pack("<II", 4, v)
        elif isinstance(v, float):
            out += struct.pack("<If", 6, v)
        else:
            out += struct.pack("<I", 8) + s(str(v))
    off = 0
    blobs = []
    for name, shape, t, data in tensors:
        data = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        out += s(name) + struct.pack("<I", len(shape))
        for d in tuple(shape)[::-1]:                 # innermost dimension first
            out += struct.pack("<Q", int(d))
        out += struct.pack("<IQ", t, off)
        blobs.append((off, data))
        off = -(-(off + data.size) // alignment) * alignment
    start = -(-len(out) // alignment) * alignment
    out += bytes(start - len(out))
    body = bytearray(off)
    for o, d in blobs:
        body[o:o + d.size] = d.tobytes()
    with open(path, "wb") as f:
        f.write(out)
        f.write(body)


def _meta_suffix(metadata, suffix, default=None):
    """`{arch}.{suffix}` for whatever architecture prefix the file uses (llama., bitnet., bitnet-b1.58., ...)."""
    arch = metadata.get("general.architecture")
    if arch and (arch + "." + suffix) in metadata:
        return metadata[arch + "." + suffix]
    for k, v in metadata.items():
        if k.endswith("." + suffix):
            return v
    return default


def load_model_tensors(reader, layout=None):
    """Every tensor the decode harness needs, by canonical name, straight from the GGUF bytes:
    returns (shape_fields, layers, emb_f16_bits, final_norm). I2_S projections stay PACKED (the bytes as they lie on
    disk, row-padded for QK256); norms and the embedding are converted to f32 / f16 bits on the host."""
    canon = {}
    for name in reader.tensors:
        c = map_tensor_name(name)
        if c is not None:
            canon[c] = name

    def need(c):
        if c not in canon:
            raise ValueError("GGUF: tensor for '%s' not found (tried the reference's aliases, weight_mapper.rs)" % c)
        return canon[c]

    def as_f32(c):
        info = reader.tensors[need(c)]
        if info.ggml_type == GGML_TYPE_F32:
            return np.frombuffer(reader.tensor_bytes(info.name, info.nelems * 4), dtype=np.float32).copy()
        if info.ggml_type == GGML_TYPE_F16:
            return np.frombuffer(reader.tensor_bytes(info.name, info.nelems * 2), dtype=np.float16).astype(np.float32)
        raise ValueError("tensor '%s': type %d where f32 / f16 is expected" % (info.name, info.ggml_type))

    emb_info = reader.tensors[need("embed_tokens.weight")]
    vocab, hidden = emb_info.shape
    if emb_info.ggml_type == GGML_TYPE_F16:
        emb = np.frombuffer(reader.tensor_bytes(emb_info.name, vocab * hidden * 2), dtype=np.uint16).copy()
    elif emb_info.ggml_type == GGML_TYPE_F32:
        emb = np.frombuffer(reader.tensor_bytes(emb_info.name, vocab * hidden * 4), dtype=np.float32).astype(np.float16).view(np.uint16)
    else:
        raise ValueError("embedding '%s': type %d, f16 or f32 expected" % (emb_info.name, emb_info.ggml_type))
    if "lm_head.weight" in canon:
        raise ValueError("GGUF: an untied lm_head ('%s') is outside the decode harness (tied embedding only)" % canon["lm_head.weight"])
    md = reader.metadata
    n_layers = int(_meta_suffix(md, "block_count", 0)) or (1 + max(int(k.split(".")[1]) for k in canon if k.startswith("layers.")))
    n_heads = int(_meta_suffix(md, "attention.head_count"))
    n_kv = int(_meta_suffix(md, "attention.head_count_kv", n_heads))
    q_rows = reader.tensors[need("layers.0.attention.q_proj.weight")].shape[0]
    ffn = reader.tensors[need("layers.0.feed_forward.gate_proj.weight")].shape[0]
    fields = dict(hidden=hidden, ffn=ffn, n_layers=n_layers, n_heads=n_heads, n_kv_heads=n_kv, head_dim=q_rows // n_heads,
                  vocab=vocab, rms_eps=float(_meta_suffix(md, "attention.layer_norm_rms_epsilon", 1e-5)),
                  rope_theta=float(_meta_suffix(md, "rope.freq_base", 500000.0)))
    keys = {"wq": "attention.q_proj", "wk": "attention.k_proj", "wv": "attention.v_proj", "wo": "attention.o_proj",
            "wgate": "feed_forward.gate_proj", "wup": "feed_forward.up_proj", "wdown": "feed_forward.down_proj"}
    layers = []
    for l in range(n_layers):
        w = {}
        for k, c in keys.items():
            name = need("layers.%d.%s.weight" % (l, c))
            info = reader.tensors[name]
            if len(info.shape) != 2:
                raise ValueError("I2_S tensor '%s' is not a matrix: shape %r" % (name, info.shape))
            rows, cols = info.shape
            flavor = reader.i2s_flavor(name)
            if flavor == I2SFlavor.GgmlQk256NoScale:
                lay, nbytes = LAYOUT_QK256, rows * (-(-cols // 256) * 64)
            elif flavor == I2SFlavor.BitNet32F16:
                lay, nbytes = LAYOUT_BITNET32_F16, rows * (cols // 32) * 10
            else:
                raise ValueError("I2_S '%s': split layout with a sibling scale tensor is not uploaded" % name)
            if layout is None:
                layout = lay
            if lay != layout:
                raise ValueError("I2_S '%s': flavour %s differs from the model's layout" % (name, flavor.name))
            data = np.zeros(nbytes, np.uint8)
            n = min(nbytes, info.size)
            data[:n] = reader.tensor_bytes(name, n)
            w[k] = data
        w["attn_norm"] = as_f32("layers.%d.attention_norm.weight" % l)
        w["ffn_norm"] = as_f32("layers.%d.post_attention_layernorm.weight" % l)
        layers.append(w)
    fields["layout"] = layout
    return fields, layers, emb, as_f32("final_norm.weight")
