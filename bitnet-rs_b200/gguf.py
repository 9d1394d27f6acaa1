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
