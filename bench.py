#!/usr/bin/env python
"""bench.py -- BitNet b1.58-2B-4T-shape batch-1 decode on B200 (BASELINE.json configs[1]).

  python bench.py --gpus N --steps K --warmup W          (N > 1: launched by torchrun)
  python bench.py --impl reference ...                   (the reference's CPU path, host cores)

A "step" is one decode step: one token per sequence through all 30 layers (RMSNorm+act-quant,
ternary q/k/v/o/gate/up/down linears, RoPE, KV append, GQA attention, SwiGLU glue), the final
norm, the tied-embedding logits and the greedy argmax. Prints ONE JSON line (rank 0).

  value     tok/s with the token fed back on device (inputs resident in HBM), CUDA events
  e2e       tok/s through the C-ABI decode call with HOST buffers: token ids H2D and the
            greedy token D2H every step, inside the timed region
  roofline  the ternary GEMVs of one step (4 fused launches per layer, 521 MB of packed
            weights > L2) replayed back to back: algorithmic bytes / measured time vs the
            measured HBM copy peak
  cpu_baseline  the CPU oracle (C++ restatement of the reference's AVX2 kernels) on a bounded
            sample, on this box's host cores
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PROMPT = 32  # default; --prompt overrides


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--layout", default="qk256", choices=["qk256", "bitnet32"])
    ap.add_argument("--prompt", type=int, default=32, help="synthetic prompt tokens decoded before timing")
    ap.add_argument("--workload", default="2b4t-decode", choices=["2b4t-decode", "8b-tp", "sweep", "prefill"],
                    help="2b4t-decode = BASELINE.json configs[1] (the judged line); 8b-tp = configs[4]: 8B-class "
                         "batch-64 decode, tensor-parallel over the launched ranks (NCCL all-reduce)")
    ap.add_argument("--tp-layers", type=int, default=0, help="8b-tp: use fewer layers (smoke runs only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=20.0)
    ap.add_argument("--tp-exchange", choices=["fused", "nccl", "both"], default="both",
                    help="8b-tp workload: how the row-parallel layers exchange (fused NVLink pulls, NCCL collectives, or both)")
    ap.add_argument("--prefill-layers", type=int, default=30)
    ap.add_argument("--no-fused-epilogue", action="store_true", help="prefill: the round-1 pipeline (stand-alone SwiGLU pass)")
    ap.add_argument("--no-bitnet32-record", action="store_true", help="skip the BitNet32-F16 decode sub-record of the default line")
    ap.add_argument("--no-gemm-record", action="store_true", help="skip the prefill GEMM sub-record of the default line")
    ap.add_argument("--no-tp", action="store_true", help="N > 1: skip the tensor-parallel sub-record of the default workload")
    return ap.parse_args()


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        if os.environ.get("BENCH_NO_CLOCKS"):   # diagnostics: is the poller itself visible in the number?
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def kernel_source_sha16():
    """Hash of the sources the persistent decode kernel is built from (ties an ncu capture to a build)."""
    import hashlib
    h = hashlib.sha256()
    for f in ("decode_mega.cu", "decode_mega.h", "common.cuh"):
        with open(os.path.join(ROOT, "bitnet-rs_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def linear_alg_bytes(shape, M):
    """W + M*K + 4*M + 4*M*N per linear call, reference-layout weight bytes."""
    qd, kd = shape.n_heads * shape.head_dim, shape.n_kv_heads * shape.head_dim
    fused = [(qd + 2 * kd, shape.hidden), (shape.hidden, qd), (2 * shape.ffn, shape.hidden),
             (shape.hidden, shape.ffn)]
    return shape.n_layers * sum(n * shape.row_bytes(k) + M * k + 4 * M + 4 * M * n for n, k in fused)


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle model (C++ restatement of the reference's kernels) on host cores
# ------------------------------------------------------------------------------------------
def cpu_decode_tok_s(shape_full, steps, warmup, budget_s):
    """tok/s of the CPU path on the FULL shape (all layers, full vocabulary), measured: the oracle model decodes
    `steps` tokens with all host threads (rows split over threads), bounded by `budget_s` of wall time, then a few
    tokens single-threaded -- which is how the reference runs this path (gemv_qk256 is a serial loop
    over rows). Returns (tok/s all threads, threads, sample text, s/step, steps timed, tok/s single thread)."""
    import bitnet_rs_b200 as b2
    from oracle import pyoracle as orc

    threads = os.cpu_count() or 1
    emb, fn = b2.synth_embedding(shape_full)
    cos_t, sin_t = b2.rope_tables(shape_full)
    m = orc.Model(shape_full, threads=threads)
    for l in range(shape_full.n_layers):
        m.set_layer(l, b2.synth_layer(shape_full, l))
    m.set_embedding(emb, fn)
    m.set_rope(cos_t, sin_t)
    t_start = time.perf_counter()
    tok = 42
    for _ in range(min(warmup, 2)):
        tok = m.step(tok)
    times = []
    for i in range(steps):
        t0 = time.perf_counter()
        tok = m.step(tok)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_start > budget_s and i >= 2:
            break
    t_full = float(np.mean(times))
    single = None
    if hasattr(m, "set_threads"):
        m.set_threads(1)
        ts = []
        for i in range(3):
            t0 = time.perf_counter()
            tok = m.step(tok)
            ts.append(time.perf_counter() - t0)
            if time.perf_counter() - t_start > 2 * budget_s:
                break
        single = 1.0 / float(np.mean(ts))
    m.close()
    sample = ("%d full-vocab decode steps of the whole %d-layer 2B-4T-shape decoder, measured (%.1f ms per step on %d "
              "threads, rows split over threads); oracle = C++ restatement of gemv_qk256_avx2 + quantize_symmetric_i8 + "
              "rms_norm" % (len(times), shape_full.n_layers, t_full * 1e3, threads))
    if single is not None:
        sample += ("; single thread, as the reference runs this path: %.2f tok/s; the reference's own published "
                   "figure for the BitNet32-F16 layout on its CI host: 0.5126 tok/s "
                   "(docs/baselines/perf/phase2_timing_i2s.md)" % single)
    return 1.0 / t_full, threads, sample, t_full, len(times), single


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import bitnet_rs_b200 as b2
    layout = b2.LAYOUT_QK256 if args.layout == "qk256" else b2.LAYOUT_BITNET32_F16
    shape = b2.ModelShape.bitnet_2b4t(max_ctx=PROMPT + 2 * (args.warmup + args.steps) + 8, layout=layout)
    steps = min(args.steps, 100)
    v, cores, sample, t_full, done, single = cpu_decode_tok_s(shape, steps, args.warmup, 120.0)
    if done < args.steps:
        sample += "; %d of the requested %d steps were timed (bounded sample)" % (done, args.steps)
    line = {
        "impl": "reference", "metric": "2B-4T decode tok/s", "value": v, "unit": "tok/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_full * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
        "config": workload_config(shape, args),
        "cpu_baseline": {"value": v, "unit": "tok/s", "cores": cores, "kind": "port", "sample": sample,
                         "single_thread_tok_s": single},
        "e2e": {"value": v, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(shape, args):
    return {
        "workload": "BitNet b1.58-2B-4T-shape full decode (30 layers, hidden 2560, FFN 6912, GQA 20/5, "
                    "vocab 128256), random-init I2_S weights, batch 1 per GPU (BASELINE.json configs[1])",
        "layout": "QK256" if shape.layout == 0 else "BitNet32-F16",
        "batch_per_gpu": shape.batch, "prompt_tokens": PROMPT,
        "parallelism": "1 GPU" if args.gpus == 1 else "%d independent replicas (one request stream per GPU)" % args.gpus,
        "l2": "inputs larger than L2: %.0f MB of packed ternary weights + 657 MB f16 LM head per step vs 126 MB L2"
              % (shape.ternary_bytes() / 1e6),
    }


def tp_measure(args, dist, world, rank, local, exchange, K, W, n_layers=None):
    """One timed run of the 8B-class batch-64 tensor-parallel decode step on the ranks of `dist` (None = 1 GPU).
    exchange: "nccl" = the all-reduces are NCCL collectives between kernels; "fused" = the consumer kernels pull
    their peers' buffers over NVLink themselves (csrc/tp_xchg.cuh). Returns a dict (same on every rank)."""
    import torch
    import bitnet_rs_b200 as b2
    from bitnet_rs_b200 import tp

    shape = b2.ModelShape.llama_8b_class(batch=64, max_ctx=PROMPT + 2 * (W + K) + 16)
    ctx = b2.Context(local)
    fabric = None
    if exchange == "fused":
        if dist is None:
            fabric = b2.TpFabric.emulated(torch, torch.device("cuda", local), 1, shape.batch, shape.hidden)[0]
        else:
            fabric = b2.TpFabric.symmetric(torch, dist, torch.device("cuda", local), shape.batch, shape.hidden)
    dec = b2.TPDecoder(ctx, shape, rank if dist is not None else 0, world if dist is not None else 1, n_layers=n_layers,
                       fabric=fabric)
    dec.tokens.copy_((torch.arange(shape.batch, device="cuda") * 7919) % shape.vocab)

    def step():
        if dist is None:
            tp.run_emulated([dec.step_gen()])
        else:
            tp.run_distributed(dec.step_gen(), dist)

    for _ in range(2):            # eager: first-use allocations (split-K scratch, NCCL channels)
        step()
    torch.cuda.synchronize()
    launches0 = b2.launch_count()
    step()
    launches_per_step = b2.launch_count() - launches0
    mode = "cuda-graph"
    graph = None
    try:
        if os.environ.get("BENCH_TP_EAGER"):
            raise RuntimeError("eager requested")
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            ctx.use_torch_stream()
            step()
        ctx.use_torch_stream()
    except Exception as ex:       # capture unavailable: time the eager loop instead
        graph, mode = None, "eager (graph capture failed: %s)" % type(ex).__name__
        ctx.use_torch_stream()
        torch.cuda.synchronize()

    def run(n):
        for _ in range(n):
            graph.replay() if graph is not None else step()

    run(W)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(K)
    e1.record()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    if dist is not None:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    nl = dec.n_layers
    out = {"tok_s": shape.batch * K / (ms * 1e-3), "ms_per_step": ms / K, "launch": mode, "clocks": clocks, "n_layers": nl,
           "launches_per_step": int(launches_per_step), "exchange": exchange,
           "collectives_per_step": (4 * nl + 1) if exchange == "nccl" else 1,
           "weight_mb_per_rank": dec.weight_bytes / 1e6, "batch": shape.batch}
    del dec, graph, fabric
    torch.cuda.empty_cache()
    return out


TP_WORKLOAD = ("8B-class ternary decoder (hidden 4096, FFN 14336, %d layers, GQA 32/8, vocab 128256), batch 64, "
               "random-init I2_S QK256 weights (BASELINE.json configs[4])")
TP_PAR = {"nccl": "tp%d: q/k/v + gate/up by rows, o + down by input columns, int32 all-reduce(SUM) + f32 all-reduce(MAX) of "
                  "the per-token absmax as NCCL collectives between the kernels",
          "fused": "tp%d: q/k/v + gate/up by rows, o + down by input columns; the int32 SUM and the absmax MAX are pulled "
                   "from the peers' symmetric-memory buffers over NVLink by the consumer kernels themselves "
                   "(sum + residual + RMSNorm + act-quant in one kernel, max + act-quant in one kernel), no collective call"}


def run_tp(args):
    """BASELINE.json configs[4]: 8B-class ternary model (hidden 4096, FFN 14336, 32 layers), batch-64
    decode, tensor-parallel over WORLD_SIZE GPUs (Megatron split). Prints one line per exchange variant."""
    import torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    K, W = args.steps, max(args.warmup, 3)
    variants = ["fused", "nccl"] if args.tp_exchange == "both" else [args.tp_exchange]
    for ex in variants:
        r = tp_measure(args, dist, world, rank, local, ex, K, W, n_layers=args.tp_layers or None)
        if rank == 0:
            line = {
                "metric": "8B-class batch-64 decode tok/s (tensor parallel)", "value": r["tok_s"],
                "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
                "config": {"workload": TP_WORKLOAD % r["n_layers"], "parallelism": TP_PAR[ex] % world, "launch": r["launch"],
                           "l2": "weights %.0f MB per rank vs 126 MB L2" % r["weight_mb_per_rank"]},
                "clocks": r["clocks"], "gpu_launches": int(r["launches_per_step"] * K),
                "launches_per_step": r["launches_per_step"], "collectives_per_step": r["collectives_per_step"],
            }
            print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def int8_library_peak(torch, sustained_iters=60):
    """cuBLASLt int8 GEMM 8192^3 (torch._int_mm) timed in THIS run: the library denominator beside the 4 500 TOPS nominal."""
    n = 8192
    a = torch.randint(-127, 128, (n, n), dtype=torch.int8, device="cuda")
    b = torch.randint(-127, 128, (n, n), dtype=torch.int8, device="cuda").t().contiguous().t()
    for _ in range(3):
        torch._int_mm(a, b)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch._int_mm(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(sustained_iters):
        torch._int_mm(a, b)
    e1.record()
    torch.cuda.synchronize()
    ops = 2.0 * n ** 3
    return ops / best / 1e9, ops / (e0.elapsed_time(e1) / sustained_iters) / 1e9


def prefill_measure(torch, b2, ctx, tokens, batch, n_layers, reps, fused=True):
    """BASELINE.json configs[3]: the ternary-linear pipeline of a 2B-4T-shape prefill (M = tokens * batch rows) --
    RMSNorm+act-quant, qkv GEMM, act-quant, o GEMM, RMSNorm+act-quant, gate/up GEMM (+ SwiGLU in its epilogue when
    `fused`), act-quant, down GEMM -- per layer, CUDA events. Attention is outside the ternary path:
    the act-quant before o reads a stand-in buffer of the attention output's shape."""
    from bitnet_rs_b200 import tp
    M = tokens * batch
    s = b2.ModelShape.bitnet_2b4t()
    H, F, QD = s.hidden, s.ffn, s.n_heads * s.head_dim
    layers = []
    for l in range(n_layers):
        w = tp.shard_layer(s, tp.fast_layer(s, l), 0, 1)          # tp = 1: the fused qkv / interleaved gate-up tensors
        ent = {k: ctx.weight(w[k][0], s.layout, w[k][1], w[k][2]) for k in ("wqkv", "wo", "wgu", "wdown")}
        ent["attn_norm"] = torch.from_numpy(w["attn_norm"]).cuda()
        ent["ffn_norm"] = torch.from_numpy(w["ffn_norm"]).cuda()
        layers.append(ent)
    x = torch.randn(M, H, device="cuda")
    att = torch.randn(M, QD, device="cuda")
    qH = torch.empty(M, H, dtype=torch.int8, device="cuda")
    qQ = torch.empty(M, QD, dtype=torch.int8, device="cuda")
    qF = torch.empty(M, F, dtype=torch.int8, device="cuda")
    sc = torch.empty(M, device="cuda")
    qkv = torch.empty(M, QD + 2 * s.n_kv_heads * s.head_dim, device="cuda")
    y = torch.empty(M, H, device="cuda")
    gu = None if fused else torch.empty(M, 2 * F, device="cuda")
    h = torch.empty(M, F, device="cuda")

    def one_pass():
        for L in layers:
            ctx.rmsnorm_actquant(x, L["attn_norm"], s.rms_eps, q=qH, scale=sc)
            ctx.linear_i8(L["wqkv"], qH, sc, y=qkv)
            ctx.actquant_i8(att, q=qQ, scale=sc)
            ctx.linear_i8(L["wo"], qQ, sc, y=y)
            ctx.rmsnorm_actquant(y, L["ffn_norm"], s.rms_eps, q=qH, scale=sc)
            if fused:
                ctx.linear_i8_swiglu(L["wgu"], qH, sc, h=h)
                ctx.actquant_i8(h, q=qF, scale=sc)
            else:
                ctx.linear_i8(L["wgu"], qH, sc, y=gu)
                h2, am = ctx.row_absmax(gu, swiglu=True)
                ctx.actquant_i8_absmax(h2, am, q=qF, scale=sc)
            ctx.linear_i8(L["wdown"], qF, sc, y=y)

    def gemms():
        for L in layers:
            ctx.linear_i8(L["wqkv"], qH, sc, y=qkv)
            ctx.linear_i8(L["wo"], qQ, sc, y=y)
            if fused:
                ctx.linear_i8_swiglu(L["wgu"], qH, sc, h=h)
            else:
                ctx.linear_i8(L["wgu"], qH, sc, y=gu)
            ctx.linear_i8(L["wdown"], qF, sc, y=y)

    def best_of(fn):
        fn()
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    l0 = b2.launch_count()
    one_pass()
    launches = b2.launch_count() - l0
    ms_all, ms_gemm = best_of(one_pass), best_of(gemms)
    # the gate/up GEMM alone (the largest; K = 2560)
    L = layers[0]
    ms_gu = best_of(lambda: [ctx.linear_i8_swiglu(L["wgu"], qH, sc, h=h) if fused else ctx.linear_i8(L["wgu"], qH, sc, y=gu) for _ in range(4)]) / 4
    per_layer = sum(r * c for r, c in s.proj_shapes().values())
    ops = 2.0 * M * per_layer * n_layers
    return {"M": M, "n_layers": n_layers, "ms_linears_and_quant": ms_all, "ms_gemm_only": ms_gemm, "int8_TOP": ops / 1e12,
            "TOPS_gemm_only": ops / ms_gemm / 1e9, "TOPS_with_quant": ops / ms_all / 1e9, "launches_per_pass": int(launches),
            "gateup_ms": ms_gu, "gateup_TOPS": 2.0 * M * 2 * F * H / ms_gu / 1e9, "fused_swiglu_epilogue": bool(fused),
            "prefill_tok_per_s_linears_only": M / (ms_all * 1e-3)}


def run_prefill(args):
    """--workload prefill: BASELINE.json configs[3], one JSON line (tensor-bound roofline)."""
    import torch
    import bitnet_rs_b200 as b2
    torch.cuda.set_device(0)
    ctx = b2.Context(0)
    sampler = ClockSampler(0)
    sampler.start()
    r = prefill_measure(torch, b2, ctx, 2048, 8, args.prefill_layers, 3, fused=not args.no_fused_epilogue)
    clocks = sampler.stop()
    burst, sustained = int8_library_peak(torch)
    line = {
        "metric": "2B-4T prefill int8 TOPS (ternary linears + act-quant)", "value": r["TOPS_with_quant"], "unit": "TOPS",
        "n_gpus": 1, "steps": 3, "warmup": 1, "ms_per_step": r["ms_linears_and_quant"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
        "config": {"workload": "2B-4T-shape prefill of a 2048-token synthetic prompt, batch 8 (M = %d token rows), %d layers, "
                               "QK256, tcgen05 kind::i8 GEMM path; linears + act-quant, attention not run "
                               "(BASELINE.json configs[3])" % (r["M"], r["n_layers"]),
                   "l2": "per layer 1.2 GB of activations / outputs stream through HBM (> 126 MB L2)"},
        "clocks": clocks, "gpu_launches": r["launches_per_pass"] * 3, "launches_per_step": r["launches_per_pass"],
        "roofline": {"bound": "tensor", "kernel": "gemm_i8_tcgen05_kernel, all four projections of %d layers" % r["n_layers"],
                     "achieved": r["TOPS_gemm_only"] / 1e3, "peak": 4.5, "unit": "PFLOP/s (int8 ops)", "frac": r["TOPS_gemm_only"] / 4500.0,
                     "peak_source": "nominal dense int8 4.5 POPS (MEASURED_PEAKS.json holds no int8 figure)",
                     "library_int8_TOPS_burst": burst, "library_int8_TOPS_sustained": sustained,
                     "frac_of_library_sustained": r["TOPS_gemm_only"] / sustained, "traffic": None},
        "prefill": r,
    }
    print(json.dumps(line), flush=True)


SWEEP_SHAPES = {"qkv": (3840, 2560), "o": (2560, 2560), "gate/up": (6912, 2560), "down": (2560, 6912)}


def _sweep_weight_bytes(rng, layout_name, rows, cols):
    if layout_name == "qk256":
        return rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
    nb = rows * cols // 32
    codes = rng.integers(0, 256, (nb, 8), dtype=np.uint8)
    sc = (rng.uniform(0.5, 1.5, nb) / np.sqrt(cols)).astype(np.float16).view(np.uint8).reshape(nb, 2)
    return np.concatenate([codes, sc], 1).ravel()


def _best_of(fn, budget_s, min_reps=3):
    fn()
    ts, t_start = [], time.perf_counter()
    while len(ts) < min_reps or (time.perf_counter() - t_start < budget_s and len(ts) < 50):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts) * 1e6


def sweep_cpu_row(layout_name, rows, cols, budget_s):
    """us per token of the reference's CPU kernel for one linear (oracle port, AVX2):
    QK256  -> gemv_qk256 (1 thread, and rows split over all host threads);
    BN32   -> KernelProvider::matmul_i2s on [k, n] unpacked code bytes, m = 1 (configs[0])."""
    from oracle import pyoracle as orc
    rng = np.random.default_rng(7)
    threads = os.cpu_count() or 1
    if layout_name == "qk256":
        qs = rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
        x = rng.standard_normal(cols).astype(np.float32)
        one = _best_of(lambda: orc.gemv_qk256(qs, x, rows, cols, cols // 4, "avx2"), budget_s)
        mt = _best_of(lambda: orc.gemv_qk256(qs, x, rows, cols, cols // 4, "mt", threads=threads), budget_s)
        return {"kernel": "gemv_qk256_avx2", "us_1thread": one, "us_all_threads": mt, "threads": threads}
    a = rng.integers(-127, 128, cols, dtype=np.int8)
    b = rng.integers(0, 3, cols * rows, dtype=np.uint8)
    one = _best_of(lambda: orc.matmul_i2s(a, b, 1, rows, cols, "avx2"), budget_s)
    return {"kernel": "matmul_i2s_avx2 (m=1, unpacked [k,n] bytes)", "us_1thread": one, "us_all_threads": None,
            "threads": 1}


def run_sweep(args):
    import torch
    import bitnet_rs_b200 as b2

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(0)
    ctx = b2.Context(0)
    peak, peak_src = measured_peak()
    i8_peak = 4500.0            # TOPS, nominal dense int8/fp8 (B200_PROFILING.md); f16 path: 2250 TFLOP/s
    ms = [1, 8, 16, 32, 64, 512, 4096]
    rows_out = []

    def graph_time(fn, iters, reps=5):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            ctx.use_torch_stream()
            for i in range(iters):
                fn(i)
        ctx.use_torch_stream()
        g.replay()
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3 / iters)
        return best

    for lname in ("qk256", "bitnet32"):
        layout = b2.LAYOUT_QK256 if lname == "qk256" else b2.LAYOUT_BITNET32_F16
        for sname, (rows, cols) in SWEEP_SHAPES.items():
            wbytes = rows * cols // 4 if lname == "qk256" else rows * cols // 32 * 10
            ncopies = min(64, max(2, (300 << 20) // wbytes + 1))
            ws = [ctx.weight(_sweep_weight_bytes(np.random.default_rng(i), lname, rows, cols), layout, rows, cols)
                  for i in range(ncopies)]
            cpu = None
            if not args.no_cpu_baseline:
                try:
                    cpu = sweep_cpu_row(lname, rows, cols, args.cpu_budget_s / 16.0)
                except Exception as ex:
                    cpu = {"failed": repr(ex)}
            for M in ms:
                q = torch.randint(-127, 128, (M, cols), dtype=torch.int8, device="cuda")
                s = torch.rand(M, device="cuda") + 0.5
                y = torch.empty(M, rows, device="cuda")
                iters = 200 if M <= 64 else (50 if M <= 512 else 20)
                l0 = b2.launch_count()
                us = graph_time(lambda i: ctx.linear_i8(ws[i % len(ws)], q, s, y=y), iters)
                alg = wbytes + M * cols + 4 * M + 4 * M * rows
                tops = 2.0 * M * rows * cols / us / 1e6
                row = {"workload": "linear_i8", "layout": lname, "shape": sname, "N": rows, "K": cols, "M": M,
                       "us": round(us, 3), "alg_bytes": alg, "GBs": round(alg / us / 1e3, 1),
                       "hbm_frac": round(alg / us / 1e3 / peak, 4), "TOPS": round(tops, 1),
                       "tensor_frac": round(tops / (i8_peak if lname == "qk256" else i8_peak / 2), 4),
                       "path": b2.last_linear_kernel(), "weight_copies": len(ws),
                       "gpu_launches": b2.launch_count() - l0}
                if M == 1 and cpu is not None:
                    row["cpu_baseline"] = cpu
                    if cpu.get("us_1thread"):
                        row["gpu_vs_cpu_1thread"] = round(cpu["us_1thread"] / us, 1)
                rows_out.append(row)
                print(json.dumps(row), flush=True)
            del ws
            torch.cuda.empty_cache()
    print(json.dumps({"summary": "sweep", "rows": len(rows_out), "hbm_peak_GBs": peak, "peak_source": peak_src,
                      "int8_peak_TOPS_nominal": i8_peak,
                      "best_hbm_frac_M1": max(r["hbm_frac"] for r in rows_out if r["M"] == 1),
                      "best_tensor_frac": max(r["tensor_frac"] for r in rows_out)}), flush=True)


def main():
    global PROMPT
    args = parse()
    PROMPT = args.prompt
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "8b-tp":
        return run_tp(args)
    if args.workload == "sweep":
        return run_sweep(args)
    if args.workload == "prefill":
        return run_prefill(args)

    import torch
    import bitnet_rs_b200 as b2

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    K, W = args.steps, max(args.warmup, 3)
    layout = b2.LAYOUT_QK256 if args.layout == "qk256" else b2.LAYOUT_BITNET32_F16
    shape = b2.ModelShape.bitnet_2b4t(max_ctx=PROMPT + 2 * (W + K) + 8, layout=layout)
    ctx = b2.Context(local)
    model = b2.Model(ctx, shape)
    model.load_synthetic()
    B = shape.batch

    # prompt: 32 synthetic token ids (i * 7919 mod vocab), one decode step each
    tok = None
    for i in range(PROMPT):
        tok = model.decode(np.array([(i * 7919 + rank) % shape.vocab] * B, np.int32))

    sampler = ClockSampler(local)
    launches0 = b2.launch_count()

    # ---- value: device-resident greedy loop, CUDA events --------------------------------
    model.decode_device(W)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    model.decode_device(K)
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1)
    launches_timed = b2.launch_count() - launches0 - W * model.launches_per_step

    # ---- e2e: the C-ABI decode call with host buffers, copies inside the timed region ----
    tok = model.get_tokens()
    for _ in range(W):
        tok = model.decode(tok)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for _ in range(K):
        tok = model.decode(tok)
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    clocks = sampler.stop()

    # ---- roofline: only the ternary GEMVs of a step, back to back ------------------------
    lps = model.launches_per_step
    ms_lin, n_lin = model.time_linears(20)
    alg = linear_alg_bytes(shape, B)
    peak, peak_src = measured_peak()
    achieved = alg / (ms_lin * 1e-3) / 1e9
    # dram__bytes_read + dram__bytes_write of the same launch from the committed ncu --set full capture
    # (profiles/r02_ncu_summary.json); not measured live -- a number taken under ncu is never a bench value. The
    # capture names the hash of the kernel sources it was taken at: a capture of other sources is refused (null).
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r02_ncu_summary.json")) as f:
            ent = json.load(f)["decode_mega_kernel_mode1_gemv_phases"]
        if n_lin == 1 and shape.layout == b2.LAYOUT_QK256 and B == 1:
            if ent.get("kernel_source_sha16") == kernel_source_sha16():
                traffic = ent["traffic_bytes_per_launch"]
                traffic_src = "profiles/r02_ncu_summary.json (ncu --set full, same launch, kernel sources %s)" % ent["kernel_source_sha16"]
            else:
                traffic_src = ("refused: profiles/r02_ncu_summary.json was captured at kernel sources %s, this build is %s"
                               % (ent.get("kernel_source_sha16"), kernel_source_sha16()))
    except Exception:
        pass

    # ---- N = 1: the same decode on the other on-disk layout (BitNet32-F16: 10-byte blocks with an inline f16 scale) ----
    bn32_rec = None
    if world == 1 and shape.layout == b2.LAYOUT_QK256 and not args.no_bitnet32_record:
        try:
            sh2 = b2.ModelShape.bitnet_2b4t(max_ctx=PROMPT + 2 * (W + K) + 8, layout=b2.LAYOUT_BITNET32_F16)
            m2 = b2.Model(ctx, sh2)
            m2.load_synthetic()
            t2 = None
            for i in range(PROMPT):
                t2 = m2.decode(np.array([(i * 7919) % sh2.vocab], np.int32))
            m2.decode_device(W)
            torch.cuda.synchronize()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record()
            m2.decode_device(K)
            f1.record()
            torch.cuda.synchronize()
            ms2 = f0.elapsed_time(f1)
            bn32_rec = {"layout": "BitNet32-F16", "tok_s": B * K / (ms2 * 1e-3), "ms_per_step": ms2 / K,
                        "launches_per_step": m2.launches_per_step, "ratio_to_qk256": (B * K / (ms2 * 1e-3)) / (B * K / (ms_dev * 1e-3)),
                        "weight_mb": sh2.ternary_bytes() / 1e6}
            m2.close()
            del m2
        except Exception as ex:
            bn32_rec = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:200])}

    # ---- N = 1: the prefill regime in short (BASELINE.json configs[3]): 2 layers' worth of the M = 16384 pipeline ----
    gemm_rec = None
    if world == 1 and not args.no_gemm_record:
        try:
            r = prefill_measure(torch, b2, ctx, 2048, 8, 2, 2, fused=True)
            gemm_rec = {"workload": "2B-4T-shape prefill, 2048 tokens x batch 8 (M = 16384), 2 of the 30 layers (identical shapes), "
                                    "tcgen05 kind::i8 GEMM path with the SwiGLU epilogue; `bench.py --workload prefill` runs all 30",
                        "TOPS_gemm_only": r["TOPS_gemm_only"], "TOPS_with_quant": r["TOPS_with_quant"],
                        "gateup_M16384_TOPS": r["gateup_TOPS"], "gateup_ms": r["gateup_ms"],
                        "frac_of_4500_TOPS_nominal": r["TOPS_gemm_only"] / 4500.0, "ms_per_layer": r["ms_linears_and_quant"] / 2}
        except Exception as ex:
            gemm_rec = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:200])}

    # ---- N > 1: the Megatron path on the same ranks (BASELINE.json configs[4]) --------------------
    tp_rec = None
    if dist is not None and not args.no_tp:
        model.close()
        del model
        torch.cuda.empty_cache()
        Kt = min(K, 16)
        try:
            fused = tp_measure(args, dist, world, rank, local, "fused", Kt, 3)
        except Exception as ex:      # symmetric memory unavailable: say so, keep the collective variant
            fused = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:200])}
        nccl = tp_measure(args, dist, world, rank, local, "nccl", Kt, 3)
        tp1 = tp_measure(args, None, 1, 0, local, "nccl", Kt, 3)          # every rank times tp = 1 on its own GPU
        best = fused if "tok_s" in fused and fused["tok_s"] >= nccl["tok_s"] else nccl
        tp_rec = {"workload": TP_WORKLOAD % best["n_layers"], "tok_s": best["tok_s"], "ms_per_step": best["ms_per_step"],
                  "exchange": best["exchange"], "parallelism": TP_PAR[best["exchange"]] % world,
                  "tp1_tok_s": tp1["tok_s"], "tp1_ms_per_step": tp1["ms_per_step"],
                  "efficiency": best["tok_s"] / (world * tp1["tok_s"]), "scaling": "strong",
                  "collectives_per_step": best["collectives_per_step"], "launches_per_step": best["launches_per_step"],
                  "launch": best["launch"], "steps": Kt,
                  "fused": {k: fused.get(k) for k in ("tok_s", "ms_per_step", "launches_per_step", "collectives_per_step", "error") if k in fused},
                  "nccl": {k: nccl[k] for k in ("tok_s", "ms_per_step", "launches_per_step", "collectives_per_step")}}

    if dist is not None:
        t = torch.tensor([ms_dev, ms_e2e], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e = float(t[0]), float(t[1])

    if rank == 0:
        line = {
            "metric": "2B-4T decode tok/s", "value": world * B * K / (ms_dev * 1e-3), "unit": "tok/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_dev / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8",
            "data": "synthetic", "config": workload_config(shape, args), "clocks": clocks,
            "e2e": {"value": world * B * K / (ms_e2e * 1e-3), "unit": "tok/s", "h2d_bytes_per_step": 4 * B,
                    "d2h_bytes_per_step": 4 * B, "ms_per_step": ms_e2e / K},
            "gpu_launches": int(launches_timed),
            "launches_per_step": lps,
            "roofline": {"bound": "hbm",
                         "kernel": "decode_mega_kernel restricted to its %d ternary GEMV phases (M=%d): one launch "
                                   "streams every packed ternary weight of the model once" % (4 * shape.n_layers, B)
                         if n_lin == 1 else "ternary GEMV launches of one step (M=%d), graph-replayed" % B,
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg / n_lin, "launches_per_pass": n_lin,
                         "us_per_launch": ms_lin * 1e3 / n_lin,
                         "pct_of_8TBs_nominal": 100.0 * achieved / 8000.0,
                         "whole_step_weight_GBs": (shape.ternary_bytes() + 2 * shape.vocab * shape.hidden)
                         / (ms_dev / K * 1e-3) / 1e9},
        }
        if tp_rec is not None:
            line["tp"] = tp_rec
        if gemm_rec is not None:
            line["gemm"] = gemm_rec
        if bn32_rec is not None:
            line["bitnet32"] = bn32_rec
        if world == 1 and not args.no_cpu_baseline:
            try:
                v, cores, sample, _, _, single = cpu_decode_tok_s(shape, 8, 1, args.cpu_budget_s)
                line["cpu_baseline"] = {"value": v, "unit": "tok/s", "cores": cores, "kind": "port", "sample": sample,
                                        "single_thread_tok_s": single}
            except Exception as ex:  # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "port",
                                        "sample": "failed: %r" % (ex,)}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
