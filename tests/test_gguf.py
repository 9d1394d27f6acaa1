"""GGUF I2_S reader + flavour detection (host logic): against the oracle's restatement of
detect_i2s_flavor / qk256_tolerance_bytes (reference vectors: formats/gguf/types.rs tests,
tests/i2s_flavor_detection.rs, i2s_close_match_priority.rs) and a synthetic GGUF v3 file."""
import os
import struct
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _s(b):
    b = b.encode()
    return struct.pack("<Q", len(b)) + b


def write_gguf(path, tensors, alignment=32, extra_kv=()):
    """tensors: list of (name, shape [rows, cols] or [n], ggml_type, bytes)."""
    kv = [("general.architecture", 8, "bitnet"), ("general.alignment", 4, alignment),
          ("bitnet.rms_norm_eps", 6, 1e-5), ("tokenizer.tokens", 9, (8, ["a", "bc"]))] + list(extra_kv)
    out = b"GGUF" + struct.pack("<IQQ", 3, len(tensors), len(kv))
    for k, t, v in kv:
        out += _s(k) + struct.pack("<I", t)
        if t == 8:
            out += _s(v)
        elif t == 4:
            out += struct.pack("<I", v)
        elif t == 6:
            out += struct.pack("<f", v)
        elif t == 9:
            et, items = v
            out += struct.pack("<IQ", et, len(items)) + b"".join(_s(i) for i in items)
    off, blobs = 0, []
    for name, shape, t, data in tensors:
        out += _s(name) + struct.pack("<I", len(shape)) + b"".join(struct.pack("<Q", d) for d in shape[::-1])
        out += struct.pack("<IQ", t, off)
        blobs.append((off, bytes(data)))
        off = -(-(off + len(data)) // alignment) * alignment
    pad = -(-len(out) // alignment) * alignment - len(out)
    out += b"\0" * pad
    body = bytearray(off)
    for o, d in blobs:
        body[o:o + len(d)] = d
    with open(path, "wb") as f:
        f.write(out + bytes(body))


def test_tolerance_and_flavor_match_the_reference(oracle, b200):
    from bitnet_rs_b200 import gguf
    for e in (0, 1, 100, 7999, 8000, 8001, 131072, 4423680, 10 ** 9):
        assert gguf.qk256_tolerance_bytes(e) == oracle.qk256_tolerance_bytes(e)
    rng = np.random.default_rng(1)
    checked = 0
    for nelems in (32, 256, 2560, 640 * 2560, 2560 * 6912, 1000, 257):
        b32, b256 = -(-nelems // 32), -(-nelems // 256)
        cands = {b32 * 8, b32 * 10, b256 * 64}
        for base in list(cands):
            for d in (-200, -9, -8, -1, 1, 8, 9, 64, 128, 200, 5000):
                cands.add(max(0, base + d))
        cands |= set(int(v) for v in rng.integers(0, b32 * 12 + 64, 20))
        for avail in sorted(cands):
            for sib in (False, True):
                for strict in (False, True):
                    try:
                        want = oracle.detect_i2s_flavor(avail, sib, nelems, strict)
                    except Exception:
                        want = None
                    if want is None:
                        with pytest.raises(ValueError, match="no valid flavor"):
                            gguf.detect_i2s_flavor(avail, sib, nelems, strict)
                    else:
                        assert gguf.detect_i2s_flavor(avail, sib, nelems, strict).name == want, (avail, sib, nelems, strict)
                    checked += 1
    assert checked > 1000
    # the reference's documented priorities: exact QK256 wins over everything
    assert gguf.detect_i2s_flavor(64, True, 256) == gguf.I2SFlavor.GgmlQk256NoScale
    assert gguf.detect_i2s_flavor(80, False, 256) == gguf.I2SFlavor.BitNet32F16


def test_reader_parses_a_synthetic_gguf(tmp_path, b200):
    from bitnet_rs_b200 import gguf
    rng = np.random.default_rng(2)
    rows, cols = 8, 512
    qk = rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
    b32 = rng.integers(0, 256, rows * cols // 32 * 10, dtype=np.uint8)
    norm = rng.standard_normal(cols).astype(np.float32)
    path = str(tmp_path / "m.gguf")
    write_gguf(path, [("blk.0.attn_q.weight", [rows, cols], 36, qk), ("blk.0.attn_norm.weight", [cols], 0, norm.tobytes()),
                      ("blk.0.ffn_down.weight", [rows, cols], 36, b32)], alignment=64)
    r = gguf.GgufReader(path)
    assert r.version == 3 and r.alignment == 64 and r.metadata["general.architecture"] == "bitnet"
    assert r.metadata["tokenizer.tokens"] == ["a", "bc"] and abs(r.metadata["bitnet.rms_norm_eps"] - 1e-5) < 1e-12
    assert [t.name for t in r.i2s_tensors()] == ["blk.0.attn_q.weight", "blk.0.ffn_down.weight"]
    assert r.tensors["blk.0.attn_q.weight"].shape == (rows, cols)
    assert r.i2s_flavor("blk.0.attn_q.weight") == gguf.I2SFlavor.GgmlQk256NoScale
    assert r.i2s_flavor("blk.0.ffn_down.weight") == gguf.I2SFlavor.BitNet32F16
    assert np.array_equal(r.tensor_bytes("blk.0.attn_q.weight", qk.size), qk)
    assert np.array_equal(r.tensor_bytes("blk.0.ffn_down.weight", b32.size), b32)
    assert np.array_equal(np.frombuffer(r.tensor_bytes("blk.0.attn_norm.weight", cols * 4), np.float32), norm)
    with pytest.raises(ValueError, match="not I2_S"):
        r.i2s_flavor("blk.0.attn_norm.weight")
    with open(str(tmp_path / "bad.gguf"), "wb") as f:
        f.write(b"GGML" + b"\0" * 64)
    with pytest.raises(ValueError, match="bad magic"):
        gguf.GgufReader(str(tmp_path / "bad.gguf"))


@pytest.mark.gpu
def test_upload_from_gguf_matches_the_oracle(tmp_path, b200, oracle):
    import torch
    from bitnet_rs_b200 import gguf
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    ctx = b200.Context(0)
    rows, cols, M = 640, 2560, 3
    qk = oracle.fill_bytes(rows * cols // 4, 0xB2000070)
    b32 = oracle.make_bitnet32(rows, cols, 0xB2000071)
    path = str(tmp_path / "w.gguf")
    write_gguf(path, [("blk.0.attn_k.weight", [rows, cols], 36, qk), ("blk.0.attn_v.weight", [rows, cols], 36, b32)])
    r = gguf.GgufReader(path)
    x = oracle.fill_normal(M * cols, 9).reshape(M, cols)
    q, s = oracle.quantize_rows_i8(x)
    w1, f1 = r.upload_i2s(ctx, "blk.0.attn_k.weight")
    w2, f2 = r.upload_i2s(ctx, "blk.0.attn_v.weight")
    assert f1 == gguf.I2SFlavor.GgmlQk256NoScale and f2 == gguf.I2SFlavor.BitNet32F16
    y1, acc1 = ctx.linear_i8(w1, torch.from_numpy(q).cuda(), torch.from_numpy(s).cuda(), want_acc=True)
    racc, ry = oracle.linear_qk256_i8(qk, rows, cols, q, s)
    assert np.array_equal(acc1.cpu().numpy(), racc) and np.array_equal(y1.cpu().numpy(), ry)
    y2 = ctx.linear_i8(w2, torch.from_numpy(q).cuda(), torch.from_numpy(s).cuda()).cpu().numpy()
    ry2 = oracle.linear_bitnet32_i8(b32, rows, cols, q, s)
    assert np.max(np.abs(y2 - ry2)) / np.max(np.abs(ry2)) < 1e-5
