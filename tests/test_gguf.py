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


# ---- whole-model load (f3): name remap + synthetic GGUF round trip ---------------------------------------
def test_tensor_name_remap_matches_weight_mapper(b200):
    # crates/bitnet-models/src/weight_mapper.rs:113-251 (vendor regexes) and :375-490 (map_tensor_name)
    from bitnet_rs_b200 import gguf
    cases = {
        "token_embd.weight": "embed_tokens.weight", "tok_embeddings.weight": "embed_tokens.weight",
        "output.weight": "lm_head.weight", "output_norm.weight": "final_norm.weight", "model.norm.weight": "final_norm.weight",
        "blk.0.attn_q.weight": "layers.0.attention.q_proj.weight", "blk.17.attn_k.weight": "layers.17.attention.k_proj.weight",
        "blk.3.attn_v.weight": "layers.3.attention.v_proj.weight", "blk.3.attn_output.weight": "layers.3.attention.o_proj.weight",
        "blk.3.attn_o.weight": "layers.3.attention.o_proj.weight", "blk.29.ffn_gate.weight": "layers.29.feed_forward.gate_proj.weight",
        "blk.29.ffn_up.weight": "layers.29.feed_forward.up_proj.weight", "blk.29.ffn_down.weight": "layers.29.feed_forward.down_proj.weight",
        "blk.5.attn_norm.weight": "layers.5.attention_norm.weight", "blk.5.ffn_norm.weight": "layers.5.post_attention_layernorm.weight",
        "blk.5.attn_sub_norm.weight": "layers.5.attention.sub_layernorm.weight",
        "model.layers.2.self_attn.q_proj.weight": "layers.2.attention.q_proj.weight",
        "layers.2.attention.wo.weight": "layers.2.attention.o_proj.weight",
        "layers.2.mlp.gate_proj.weight": "layers.2.feed_forward.gate_proj.weight",
        "layers.2.feed_forward.w2.weight": "layers.2.feed_forward.down_proj.weight",
        "model.layers.9.input_layernorm.weight": "layers.9.attention_norm.weight",
        "layers.9.ffn_norm.weight": "layers.9.post_attention_layernorm.weight",
        "blk.1.unknown.weight": None, "rope_freqs.weight": None,
    }
    for k, v in cases.items():
        assert gguf.map_tensor_name(k) == v, k
    assert gguf.qk256_key("blk.0.attn_q.weight") == "blk.0.attn_q.weight.qk256_qs"       # loader.rs:1507
    assert gguf.expected_qk256_shape(2560, 2560) == (2560, 640)                            # qk256_utils.rs:19-56
    assert gguf.expected_qk256_shape(640, 300) == (640, 128)


def _synthetic_gguf(b200, shape, path, emb_type=1):
    """The model `Model.load_synthetic` builds, written as a GGUF with llama.cpp tensor names."""
    from bitnet_rs_b200 import gguf
    names = {"wq": "attn_q", "wk": "attn_k", "wv": "attn_v", "wo": "attn_output", "wgate": "ffn_gate", "wup": "ffn_up",
             "wdown": "ffn_down"}
    tensors = []
    emb, fn = b200.synth_embedding(shape)
    tensors.append(("token_embd.weight", (shape.vocab, shape.hidden), gguf.GGML_TYPE_F16, emb))
    for l in range(shape.n_layers):
        w = b200.synth_layer(shape, l)
        for k, (r, c) in shape.proj_shapes().items():
            tensors.append(("blk.%d.%s.weight" % (l, names[k]), (r, c), gguf.GGML_TYPE_I2_S, w[k]))
        tensors.append(("blk.%d.attn_norm.weight" % l, (shape.hidden,), gguf.GGML_TYPE_F32, w["attn_norm"]))
        tensors.append(("blk.%d.ffn_norm.weight" % l, (shape.hidden,), gguf.GGML_TYPE_F32, w["ffn_norm"]))
    tensors.append(("output_norm.weight", (shape.hidden,), gguf.GGML_TYPE_F32, fn))
    md = {"general.architecture": "bitnet-b1.58", "bitnet-b1.58.block_count": shape.n_layers,
          "bitnet-b1.58.embedding_length": shape.hidden, "bitnet-b1.58.feed_forward_length": shape.ffn,
          "bitnet-b1.58.attention.head_count": shape.n_heads, "bitnet-b1.58.attention.head_count_kv": shape.n_kv_heads,
          "bitnet-b1.58.attention.layer_norm_rms_epsilon": float(shape.rms_eps), "bitnet-b1.58.rope.freq_base": float(shape.rope_theta)}
    gguf.write_gguf(path, md, tensors)


@pytest.mark.parametrize("layout_name", ["LAYOUT_QK256", "LAYOUT_BITNET32_F16"])
def test_load_model_tensors_from_synthetic_gguf(b200, tmp_path, layout_name):
    # host part of the whole-model load: shape from the metadata, every tensor found through the remap, bytes intact
    from bitnet_rs_b200 import gguf
    layout = getattr(b200, layout_name)
    shape = b200.ModelShape(hidden=256, ffn=512, n_layers=2, n_heads=4, n_kv_heads=2, head_dim=64, vocab=300, max_ctx=16, layout=layout)
    p = str(tmp_path / "m.gguf")
    _synthetic_gguf(b200, shape, p)
    r = gguf.GgufReader(p)
    fields, layers, emb, fn = gguf.load_model_tensors(r)
    r.close()
    for k in ("hidden", "ffn", "n_layers", "n_heads", "n_kv_heads", "head_dim", "vocab", "layout"):
        assert fields[k] == getattr(shape, k), k
    assert abs(fields["rope_theta"] - shape.rope_theta) < 1e-3 and abs(fields["rms_eps"] - shape.rms_eps) < 1e-9
    for l in range(2):
        w = b200.synth_layer(shape, l)
        for k in w:
            assert np.array_equal(layers[l][k], w[k]), (l, k)
    e0, f0 = b200.synth_embedding(shape)
    assert np.array_equal(emb, e0) and np.array_equal(fn, f0)


@pytest.mark.gpu
@pytest.mark.parametrize("layout_name", ["LAYOUT_QK256", "LAYOUT_BITNET32_F16"])
def test_model_load_gguf_decodes_like_load_synthetic(b200, tmp_path, layout_name):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    layout = getattr(b200, layout_name)
    shape = b200.ModelShape(hidden=256, ffn=512, n_layers=2, n_heads=4, n_kv_heads=2, head_dim=64, vocab=512, max_ctx=32, layout=layout)
    p = str(tmp_path / "m.gguf")
    _synthetic_gguf(b200, shape, p)
    ctx = b200.Context(0)
    a = b200.Model(ctx, shape)
    a.load_synthetic()
    b = b200.Model.load_gguf(ctx, p, max_ctx=32)
    assert b.shape.hidden == 256 and b.shape.n_kv_heads == 2 and b.shape.layout == layout
    ta = tb = np.array([7], np.int32)
    for _ in range(8):
        ta, la = a.decode(ta, want_logits=True)
        tb, lb = b.decode(tb, want_logits=True)
        assert np.array_equal(ta, tb) and np.array_equal(la, lb)
    a.close()
    b.close()
