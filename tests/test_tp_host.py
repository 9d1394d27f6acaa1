"""Tensor-parallel host logic on CPU (no GPU): shard byte ranges, divisibility rules, and the
row-parallel exchange protocol (all-reduce MAX of the per-token absmax, all-reduce SUM of int32
partials) run over torch.distributed `gloo` with world_size 2, checked bit for bit against the
unsharded oracle. The oracle stands in for the CUDA kernels here (test infrastructure only)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_shard_rows_and_cols_reassemble(b200):
    from bitnet_rs_b200 import tp
    rng = np.random.default_rng(0)
    for layout, blk in ((b200.LAYOUT_QK256, 256), (b200.LAYOUT_BITNET32_F16, 32)):
        rows, cols = 12, 1024
        rb = tp.row_bytes(layout, cols)
        data = rng.integers(0, 256, rows * rb, dtype=np.uint8)
        parts = [tp.shard_rows(data, layout, rows, cols, r, r + 3) for r in range(0, rows, 3)]
        assert np.array_equal(np.concatenate(parts), data)
        cparts = [tp.shard_cols(data, layout, rows, cols, c, c + 256).reshape(rows, -1) for c in range(0, cols, 256)]
        assert np.array_equal(np.concatenate(cparts, axis=1).reshape(-1), data)
        with pytest.raises(ValueError, match="splits"):
            tp.shard_cols(data, layout, rows, cols, blk // 2, cols)


def test_divisibility_rules(b200):
    from bitnet_rs_b200 import tp
    s8 = b200.ModelShape.llama_8b_class()
    for n in (1, 2, 4, 8):
        tp.check_divisible(s8, n)                       # 8B-class shards at tp 2/4/8
    s2 = b200.ModelShape.bitnet_2b4t()
    with pytest.raises(ValueError):                     # 5 KV heads: 2B-4T runs as replicas
        tp.check_divisible(s2, 2)
    sh = tp.shard_layer(s8, {k: np.zeros(r * tp.row_bytes(s8.layout, c), np.uint8)
                             for k, (r, c) in s8.proj_shapes().items()} | {"attn_norm": 0, "ffn_norm": 0}, 3, 8)
    assert sh["wqkv"][1:] == (6 * 128, 4096) and sh["wo"][1:] == (4096, 512)
    assert sh["wgu"][1:] == (2 * 1792, 4096) and sh["wdown"][1:] == (4096, 1792)


def _quant_with_absmax(orc, x, absmax):
    """quantize_symmetric_i8 of each row with a given absmax: append the absmax as one more element."""
    aug = np.concatenate([x, absmax[:, None].astype(np.float32)], axis=1)
    q, s = orc.quantize_rows_i8(aug)
    return np.ascontiguousarray(q[:, :-1]), s


def _rank_main(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sys.path.insert(0, ROOT)
        from oracle import pyoracle as orc
        import bitnet_rs_b200 as b2
        from bitnet_rs_b200 import tp
        rows, cols, M = 96, 1024, 5
        qs = orc.fill_bytes(rows * cols // 4, 0xB2000010)
        x = orc.fill_normal(M * cols, 77, 2.0).reshape(M, cols)
        x[3] = 0.0                                                   # an all-zero token: scale 0 on every rank
        kl = cols // world
        xs = np.ascontiguousarray(x[:, rank * kl:(rank + 1) * kl])
        am = torch.from_numpy(np.abs(xs).max(axis=1).astype(np.float32))
        dist.all_reduce(am, op=dist.ReduceOp.MAX)                    # exchange 1: global per-token absmax
        q, s = _quant_with_absmax(orc, xs, am.numpy())
        wl = tp.shard_cols(qs, b2.LAYOUT_QK256, rows, cols, rank * kl, (rank + 1) * kl)
        acc, _ = orc.linear_qk256_i8(wl, rows, kl, q, s)
        acc_t = torch.from_numpy(acc.astype(np.int32))
        dist.all_reduce(acc_t, op=dist.ReduceOp.SUM)                 # exchange 2: int32 partials
        qf, sf = orc.quantize_rows_i8(x)
        racc, ry = orc.linear_qk256_i8(qs, rows, cols, qf, sf)
        ok = bool(np.array_equal(acc_t.numpy(), racc)) and bool(np.array_equal(s, sf))
        y = s[:, None] * acc_t.numpy().astype(np.float32)
        ok = ok and bool(np.array_equal(y.astype(np.float32), ry))
        out[rank] = ok
    finally:
        dist.destroy_process_group()


def test_row_parallel_protocol_gloo_world2(oracle):
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_rank_main, args=(2, port, out), nprocs=2, join=True)
    assert out.get(0) is True and out.get(1) is True
