"""Tensor-parallel decode over the ternary linears (BASELINE.json configs[4]).

One process per GPU. Each layer's weights are split Megatron-style: q/k/v and gate/up by output
rows (whole heads / FFN columns, no exchange), o and down by input columns with an all-reduce after
them. The reference only simulates these semantics in-process
(crates/bitnet-inference/src/tensor_parallel.rs:123-290: partition_tensor, all_reduce Sum / Max).

Bit-exactness across tp: the row-parallel layers quantise their K/tp activation slice with the
GLOBAL per-token absmax (all-reduce MAX of M floats), run the packed matmul into int32 partial
accumulators and all-reduce those as int32 (SUM) -- integer addition, so the reduced accumulator
equals the tp = 1 accumulator bit for bit, and the scale/residual epilogue is the same multiply.

torch.distributed (NCCL over NVLink) is the plumbing for the two exchanges; every kernel between
them is from libbitnet_b200.so. The decoder step is written as a generator that yields at each
exchange, so the same code runs (a) under torch.distributed, one rank per GPU, and (b) emulated:
several ranks' shards on ONE GPU stepped in lockstep with the reductions done in place -- that is
how the single-GPU test suite checks tp = 2/4 against tp = 1.

The LM head here is a plain f16 library GEMM on a vocabulary shard (torch.matmul, cuBLAS) followed
by an all-gather of per-rank (max, argmax): it is outside the ternary path.
"""
import numpy as np

from .device import LAYOUT_QK256
from .model import rope_tables, synth_embedding, synth_uniform


# --------------------------------------------------------------------------------------------
# host-side sharding of packed tensors (numpy, no compute): byte ranges in the reference layouts
# --------------------------------------------------------------------------------------------
def row_bytes(layout, cols):
    return -(-cols // 256) * 64 if layout == LAYOUT_QK256 else cols // 32 * 10


def shard_rows(data, layout, rows, cols, r0, r1):
    """Rows [r0, r1) of a packed [rows, cols] tensor: one contiguous byte range in both layouts."""
    rb = row_bytes(layout, cols)
    a = np.asarray(data, dtype=np.uint8).reshape(rows, rb)
    return np.ascontiguousarray(a[r0:r1]).reshape(-1)


def shard_cols(data, layout, rows, cols, c0, c1):
    """Columns [c0, c1) of every row: a strided gather; c0, c1 must sit on block boundaries
    (256 for QK256, 32 for BitNet32-F16) so that no block is split."""
    blk, bb = (256, 64) if layout == LAYOUT_QK256 else (32, 10)
    if c0 % blk or c1 % blk:
        raise ValueError("column shard [%d, %d) splits a %d-element block" % (c0, c1, blk))
    rb = row_bytes(layout, cols)
    a = np.asarray(data, dtype=np.uint8).reshape(rows, rb)
    return np.ascontiguousarray(a[:, c0 // blk * bb:c1 // blk * bb]).reshape(-1)


def interleave_rows(a, b, layout, rows, cols):
    """Rows a0, b0, a1, b1, ... (gate/up pairs, so the SwiGLU glue reads adjacent outputs)."""
    rb = row_bytes(layout, cols)
    a2 = np.asarray(a, dtype=np.uint8).reshape(rows, rb)
    b2 = np.asarray(b, dtype=np.uint8).reshape(rows, rb)
    return np.ascontiguousarray(np.stack([a2, b2], axis=1)).reshape(-1)


def check_divisible(shape, tp):
    """whole heads per rank, and the row-parallel K/tp on a block boundary."""
    blk = 256 if shape.layout == LAYOUT_QK256 else 32
    qd = shape.n_heads * shape.head_dim
    if shape.n_heads % tp or shape.n_kv_heads % tp:
        raise ValueError("tp=%d does not divide heads %d / kv heads %d" % (tp, shape.n_heads, shape.n_kv_heads))
    if (qd // tp) % blk or shape.ffn % tp or (shape.ffn // tp) % blk:
        raise ValueError("tp=%d: o/down input slices (%d, %d) are not multiples of the %d-element block"
                         % (tp, qd // tp, shape.ffn // tp, blk))
    if shape.vocab % tp:
        raise ValueError("tp=%d does not divide the vocabulary %d" % (tp, shape.vocab))


def fast_layer(shape, layer, seed=0x7B200000):
    """Random-init layer for the TP configs (numpy PCG64 per tensor: the 8B-class model has 7 G weights)."""
    out = {}
    for i, (name, (r, c)) in enumerate(shape.proj_shapes().items()):
        rng = np.random.default_rng([seed, layer, i])
        out[name] = rng.integers(0, 256, r * row_bytes(shape.layout, c), dtype=np.uint8)
    out["attn_norm"] = synth_uniform(shape.hidden, seed + layer * 16 + 8, 0.5, 1.5)
    out["ffn_norm"] = synth_uniform(shape.hidden, seed + layer * 16 + 9, 0.5, 1.5)
    return out


def shard_layer(shape, w, rank, tp):
    """The four fused local tensors of one layer for `rank` of `tp` (host bytes + local shapes)."""
    L, H, F, D = shape.layout, shape.hidden, shape.ffn, shape.head_dim
    nh, nkv = shape.n_heads // tp, shape.n_kv_heads // tp
    qd, kd = shape.n_heads * D, shape.n_kv_heads * D
    q = shard_rows(w["wq"], L, qd, H, rank * nh * D, (rank + 1) * nh * D)
    k = shard_rows(w["wk"], L, kd, H, rank * nkv * D, (rank + 1) * nkv * D)
    v = shard_rows(w["wv"], L, kd, H, rank * nkv * D, (rank + 1) * nkv * D)
    fl = F // tp
    g = shard_rows(w["wgate"], L, F, H, rank * fl, (rank + 1) * fl)
    u = shard_rows(w["wup"], L, F, H, rank * fl, (rank + 1) * fl)
    return {
        "wqkv": (np.concatenate([q, k, v]), (nh + 2 * nkv) * D, H),
        "wo": (shard_cols(w["wo"], L, H, qd, rank * nh * D, (rank + 1) * nh * D), H, nh * D),
        "wgu": (interleave_rows(g, u, L, fl, H), 2 * fl, H),
        "wdown": (shard_cols(w["wdown"], L, H, F, rank * fl, (rank + 1) * fl), H, fl),
        "attn_norm": w["attn_norm"], "ffn_norm": w["ffn_norm"],
    }


# --------------------------------------------------------------------------------------------
class TpFabric:
    """The peer-visible buffers of the fused exchanges (csrc/tp_xchg.cuh) for one rank and the pointers to every
    rank's copy. Per rank, one block of `nbytes` laid out as
        acc_o [M, H] i32 | acc_down [M, H] i32 | amax_o [M] f32 | amax_down [M] f32 | flags [4][8] u32
    `local` is this rank's block as a uint8 tensor, `bases[r]` the address of rank r's block on THIS device."""
    SLOT_O_MAX, SLOT_O_SUM, SLOT_DOWN_MAX, SLOT_DOWN_SUM = 0, 1, 2, 3

    @staticmethod
    def nbytes(M, H):
        return 2 * M * H * 4 + 2 * M * 4 + 4 * 8 * 4

    def __init__(self, torch, local, bases, rank, M, H):
        self.torch, self.local, self.bases, self.rank, self.n = torch, local, list(bases), rank, len(bases)
        self.M, self.H = M, H
        o = 0
        self.off = {}
        for name, n in (("acc_o", M * H * 4), ("acc_down", M * H * 4), ("amax_o", M * 4), ("amax_down", M * 4), ("flags", 128)):
            self.off[name] = o
            o += n
        self.acc_o = local[self.off["acc_o"]:self.off["acc_o"] + M * H * 4].view(torch.int32).view(M, H)
        self.acc_down = local[self.off["acc_down"]:self.off["acc_down"] + M * H * 4].view(torch.int32).view(M, H)
        self.amax_o = local[self.off["amax_o"]:self.off["amax_o"] + M * 4].view(torch.float32)
        self.amax_down = local[self.off["amax_down"]:self.off["amax_down"] + M * 4].view(torch.float32)
        self.counters = torch.zeros(8, dtype=torch.int32, device=local.device)   # local, not peer-visible
        self._data_of = {self.SLOT_O_MAX: "amax_o", self.SLOT_O_SUM: "acc_o", self.SLOT_DOWN_MAX: "amax_down",
                         self.SLOT_DOWN_SUM: "acc_down"}

    @classmethod
    def emulated(cls, torch, device, n_ranks, M, H):
        """n_ranks fabrics on ONE device (tests): plain device tensors, every 'peer' pointer is local."""
        blocks = [torch.zeros(cls.nbytes(M, H), dtype=torch.uint8, device=device) for _ in range(n_ranks)]
        bases = [b.data_ptr() for b in blocks]
        return [cls(torch, blocks[r], bases, r, M, H) for r in range(n_ranks)]

    @classmethod
    def symmetric(cls, torch, dist, device, M, H):
        """One rank per GPU: torch symmetric memory (P2P-mapped over NVLink) is the plumbing for the addresses."""
        import torch.distributed._symmetric_memory as symm_mem
        local = symm_mem.empty(cls.nbytes(M, H), dtype=torch.uint8, device=device)
        local.zero_()
        hdl = symm_mem.rendezvous(local, dist.group.WORLD)
        torch.cuda.synchronize()
        dist.barrier()
        fab = cls(torch, local, [int(p) for p in hdl.buffer_ptrs], dist.get_rank(), M, H)
        fab._hdl = hdl
        return fab

    def desc(self, slot, phase):
        from . import ffi
        xc = ffi.TpXchg()
        for r in range(self.n):
            xc.peer_data[r] = self.bases[r] + self.off[self._data_of[slot]]
            xc.peer_flags[r] = self.bases[r] + self.off["flags"]
        xc.counters = self.counters.data_ptr()
        xc.n_ranks, xc.rank, xc.slot, xc.phase = self.n, self.rank, slot, phase
        return xc


# --------------------------------------------------------------------------------------------
class TPDecoder:
    """One rank's shard of a ternary decoder; `step_gen` yields ("max" | "sum" | "gather", tensor)
    at every exchange and expects the tensor to hold the reduced value when resumed."""

    def __init__(self, ctx, shape, rank=0, tp=1, layer_fn=fast_layer, n_layers=None, fabric=None, emulated=False):
        """fabric: a TpFabric -> the exchanges are pulled by the consumer kernels themselves (no collective calls);
        emulated: several ranks share one GPU, so every exchange kernel is split into a signal launch and a work
        launch with a lockstep point between them."""
        import torch
        if shape.layout != LAYOUT_QK256:
            raise ValueError("the TP decoder covers the QK256 layout (int32 partials reduce exactly)")
        check_divisible(shape, tp)
        self.torch, self.ctx, self.shape, self.rank, self.tp = torch, ctx, shape, rank, tp
        self.n_layers = shape.n_layers if n_layers is None else n_layers
        dev = torch.device("cuda", ctx.device)
        self.dev = dev
        self.layers = []
        for l in range(self.n_layers):
            sh = shard_layer(shape, layer_fn(shape, l), rank, tp)
            ent = {k: ctx.weight(sh[k][0], shape.layout, sh[k][1], sh[k][2]) for k in ("wqkv", "wo", "wgu", "wdown")}
            ent["attn_norm"] = torch.from_numpy(np.ascontiguousarray(sh["attn_norm"])).to(dev)
            ent["ffn_norm"] = torch.from_numpy(np.ascontiguousarray(sh["ffn_norm"])).to(dev)
            self.layers.append(ent)
        emb, fn = synth_embedding(shape)
        self.emb = torch.from_numpy(emb.reshape(shape.vocab, shape.hidden).view(np.float16)).to(dev)
        self.final_norm = torch.from_numpy(fn).to(dev)
        cos_t, sin_t = rope_tables(shape)
        self.cos, self.sin = torch.from_numpy(cos_t).to(dev), torch.from_numpy(sin_t).to(dev)
        B, D = shape.batch, shape.head_dim
        self.nh, self.nkv = shape.n_heads // tp, shape.n_kv_heads // tp
        self.kc = torch.zeros((self.n_layers, B, self.nkv, shape.max_ctx, D), dtype=torch.float32, device=dev)
        self.vc = torch.zeros_like(self.kc)
        self.pos = torch.zeros((), dtype=torch.int32, device=dev)
        self.tokens = torch.zeros((B,), dtype=torch.int64, device=dev)
        self.v0 = rank * (shape.vocab // tp)
        self.emb_shard = self.emb[self.v0:self.v0 + shape.vocab // tp]
        self.weight_bytes = sum(e[k].nbytes for e in self.layers for k in ("wqkv", "wo", "wgu", "wdown"))
        self._pos_host = 0   # host mirror of self.pos: the kernels index the caches and RoPE tables with it unchecked
        self.fabric, self.emulated = fabric, emulated

    def reset(self):
        self.pos.zero_()
        self._pos_host = 0

    def layers_gen(self, x):
        """The ternary part of one step on hidden state x [B, hidden] (updated in place)."""
        ctx, s = self.ctx, self.shape
        for l, L in enumerate(self.layers):
            q8, s8 = ctx.rmsnorm_actquant(x, L["attn_norm"], s.rms_eps)          # replicated, full hidden
            qkv = ctx.linear_i8(L["wqkv"], q8, s8)                                 # column-parallel: local heads
            att = ctx.rope_attn_decode(qkv, self.cos, self.sin, self.kc[l], self.vc[l], self.pos, self.nh, self.nkv,
                                       s.head_dim, s.max_ctx)
            am = ctx.row_absmax(att)
            yield "max", am                                                        # global per-token absmax
            q8o, so = ctx.actquant_i8_absmax(att, am)
            _, acc = ctx.linear_i8(L["wo"], q8o, so, want_acc=True)                # row-parallel: int32 partials
            yield "sum", acc
            ctx.residual_acc(acc, so, x)
            q8, s8 = ctx.rmsnorm_actquant(x, L["ffn_norm"], s.rms_eps)
            gu = ctx.linear_i8(L["wgu"], q8, s8)                                   # column-parallel, (gate, up) interleaved
            h, am = ctx.row_absmax(gu, swiglu=True)
            yield "max", am
            q8d, sd = ctx.actquant_i8_absmax(h, am)
            _, acc = ctx.linear_i8(L["wdown"], q8d, sd, want_acc=True)
            yield "sum", acc
            ctx.residual_acc(acc, sd, x)

    def _xchg(self, slot, fn):
        """Run exchange kernel fn(desc) for `slot`: one launch on a real rank; signal, lockstep point, work when emulated."""
        if self.emulated:
            fn(self.fabric.desc(slot, 1))
            yield "sync", None
            return fn(self.fabric.desc(slot, 2))
        return fn(self.fabric.desc(slot, 3))
        yield  # pragma: no cover (makes this a generator in both branches)

    def layers_fused_gen(self, x):
        """layers_gen with the four exchanges of a layer fused into their consumers (csrc/tp_xchg.cuh): 11 launches per
        layer, no collective call. Returns the final-norm output (the last exchange applies the final norm)."""
        ctx, s, fab = self.ctx, self.shape, self.fabric
        q8, s8 = ctx.rmsnorm_actquant(x, self.layers[0]["attn_norm"], s.rms_eps)
        xn = None
        for l, L in enumerate(self.layers):
            qkv = ctx.linear_i8(L["wqkv"], q8, s8)
            att = ctx.rope_attn_decode(qkv, self.cos, self.sin, self.kc[l], self.vc[l], self.pos, self.nh, self.nkv,
                                       s.head_dim, s.max_ctx)
            ctx.row_absmax(att, absmax=fab.amax_o)
            q8o, so = yield from self._xchg(fab.SLOT_O_MAX, lambda d: ctx.tp_xchg_max_actquant(d, att))
            ctx.linear_i8(L["wo"], q8o, so, acc_out=fab.acc_o)
            q8, s8 = yield from self._xchg(fab.SLOT_O_SUM, lambda d: ctx.tp_xchg_sum_rmsnorm_actquant(
                d, so, x, L["ffn_norm"], s.rms_eps))
            gu = ctx.linear_i8(L["wgu"], q8, s8)
            h, _ = ctx.row_absmax(gu, swiglu=True, absmax=fab.amax_down)
            q8d, sd = yield from self._xchg(fab.SLOT_DOWN_MAX, lambda d: ctx.tp_xchg_max_actquant(d, h))
            ctx.linear_i8(L["wdown"], q8d, sd, acc_out=fab.acc_down)
            last = l + 1 == len(self.layers)
            gamma = self.final_norm if last else self.layers[l + 1]["attn_norm"]
            r = yield from self._xchg(fab.SLOT_DOWN_SUM, lambda d: ctx.tp_xchg_sum_rmsnorm_actquant(
                d, sd, x, gamma, s.rms_eps, want_normed=last))
            if last:
                xn = r[2]
            else:
                q8, s8 = r
        return xn

    def step_gen(self):
        """One greedy decode step: self.tokens [B] -> self.tokens, self.pos += 1."""
        torch, s = self.torch, self.shape
        if self._pos_host >= s.max_ctx:
            raise ValueError("TPDecoder: context full (position %d, max_ctx %d): call reset()" % (self._pos_host, s.max_ctx))
        self._pos_host += 1
        x = self.emb[self.tokens].float()
        if self.fabric is not None:
            xn = yield from self.layers_fused_gen(x)
        else:
            yield from self.layers_gen(x)
            _, _, xn = self.ctx.rmsnorm_actquant(x, self.final_norm, s.rms_eps, want_normed=True)
        # LM head on the vocabulary shard: a library f16 GEMM (outside the ternary path). The f32 row is
        # fed as hi + lo f16 halves (xn = hi + lo to ~2^-22), so the logits are those of f32 activations x f16 weights
        # up to fp32 summation order -- not the f16-rounded activations of round 1 -- and ties go to the LOWEST index,
        # as the device argmax of the single-GPU path does (torch.max / argmax do not promise that).
        hi = xn.half()
        lo = (xn - hi.float()).half()
        try:     # f16 x f16 with an f32 result (aten::mm.dtype, CUDA): the f16 output rounding would undo the split
            both = torch.mm(torch.cat([hi, lo], dim=0), self.emb_shard.t(), out_dtype=torch.float32)
        except (TypeError, RuntimeError, NotImplementedError):
            both = torch.matmul(torch.cat([hi, lo], dim=0), self.emb_shard.t()).float()
        logits = both[:s.batch] + both[s.batch:]
        val = logits.max(dim=1).values
        big = torch.iinfo(torch.int64).max
        ar = torch.arange(logits.shape[1], device=self.dev)
        idx = torch.where(logits == val[:, None], ar[None, :], torch.full_like(ar, big)[None, :]).min(dim=1).values
        pair = torch.stack([val, (idx + self.v0).float()], dim=1).contiguous()    # [B, 2]
        gathered = torch.empty((self.tp, s.batch, 2), dtype=torch.float32, device=self.dev)
        gathered[self.rank] = pair
        yield "gather", gathered                                                    # all ranks' (max, argmax)
        gv, gi = gathered[:, :, 0], gathered[:, :, 1]
        best = gv.max(dim=0).values
        self.tokens.copy_(torch.where(gv == best[None, :], gi, torch.full_like(gi, float(s.vocab))).min(dim=0).values.long())
        self.pos += 1
        self.last_hidden = x


def run_distributed(gen, dist, group=None):
    """Drive one rank's generator with torch.distributed collectives (NCCL) on the current stream."""
    for op, t in gen:
        if op == "sync":
            continue
        if op == "max":
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        elif op == "sum":
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        else:
            dist.all_gather_into_tensor(t.view(-1), t[dist.get_rank(group)].clone().view(-1), group=group)


def run_emulated(gens):
    """Step several ranks' generators in lockstep on one device; reductions written back in place."""
    import torch
    gens = list(gens)
    while True:
        ys = []
        for g in gens:
            try:
                ys.append(next(g))
            except StopIteration:
                ys.append(None)
        if all(y is None for y in ys):
            return
        if any(y is None for y in ys):
            raise RuntimeError("ranks fell out of lockstep")
        op = ys[0][0]
        ts = [y[1] for y in ys]
        if op == "sync":
            continue
        if op == "max":
            r = torch.stack(ts).amax(dim=0)
            for t in ts:
                t.copy_(r)
        elif op == "sum":
            r = torch.stack(ts).sum(dim=0, dtype=torch.int32)
            for t in ts:
                t.copy_(r)
        else:
            full = ts[0].clone()
            for k, t in enumerate(ts):
                full[k] = t[k]
            for t in ts:
                t.copy_(full)
