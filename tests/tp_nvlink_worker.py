"""Worker of tests/test_gpu_tp_nvlink.py: one process per GPU (torchrun). Every rank runs its shard of a small GQA decoder
with the fused NVLink exchanges (torch symmetric memory), then with NCCL collectives, eagerly and from a CUDA graph; rank 0
also runs the unsharded tp = 1 decoder. Hidden states and tokens must be bit-identical in all of them."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2  # noqa: E402
from bitnet_rs_b200 import tp  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # divisible down to tp = 8: 16 / 8 heads, and 2048 / 8 = 256 input columns per rank for o and down (one QK256 block)
    shape = b2.ModelShape(hidden=2048, ffn=2048, n_layers=3, n_heads=16, n_kv_heads=8, head_dim=128, vocab=4096, max_ctx=32,
                          batch=16)
    ctx = b2.Context(local)
    tok0 = (torch.arange(shape.batch, device="cuda") * 7919) % shape.vocab
    steps = 6

    def run(dec, n, graph=False):
        dec.reset()
        dec.tokens.copy_(tok0)
        hs, ts = [], []
        g = None
        for i in range(n):
            if graph and i == 2:       # two eager steps first (first-use allocations, kernel selection), then capture
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    ctx.use_torch_stream()
                    tp.run_distributed(dec.step_gen(), dist)     # records the step (host position + 1), runs nothing
                ctx.use_torch_stream()
                g.replay()
            elif g is not None:
                g.replay()
                dec._pos_host += 1     # replays advance the device-side position only
            else:
                tp.run_distributed(dec.step_gen(), dist)
            torch.cuda.synchronize()
            hs.append(dec.last_hidden.clone())
            ts.append(dec.tokens.clone())
        return hs, ts

    try:
        fabric = b2.TpFabric.symmetric(torch, dist, torch.device("cuda", local), shape.batch, shape.hidden)
        avail = torch.tensor([1], device="cuda")
    except Exception as ex:      # no P2P / symmetric memory between these GPUs: nothing to test here
        fabric, avail = None, torch.tensor([0], device="cuda")
        why = repr(ex)
    dist.all_reduce(avail, op=dist.ReduceOp.MIN)
    if int(avail[0]) == 0:
        if rank == 0:
            print("TP_NVLINK_UNAVAILABLE: torch symmetric memory could not be set up (%s)" % (why if fabric is None else "on a peer"))
        dist.barrier()
        dist.destroy_process_group()
        sys.exit(0)
    fused = b2.TPDecoder(ctx, shape, rank, world, fabric=fabric)
    h_f, t_f = run(fused, steps)
    h_g, t_g = run(fused, steps, graph=True)
    nccl = b2.TPDecoder(ctx, shape, rank, world)
    h_n, t_n = run(nccl, steps)
    ok = all(torch.equal(a, b) for a, b in zip(h_f, h_n)) and all(torch.equal(a, b) for a, b in zip(t_f, t_n))
    ok_g = all(torch.equal(a, b) for a, b in zip(t_f, t_g)) and all(torch.equal(a, b) for a, b in zip(h_f, h_g))
    ok_1 = True
    if rank == 0:
        ref = b2.TPDecoder(ctx, shape, 0, 1)
        ref.tokens.copy_(tok0)
        for i in range(steps):
            tp.run_emulated([ref.step_gen()])
            ok_1 = ok_1 and torch.equal(ref.last_hidden, h_f[i]) and torch.equal(ref.tokens, t_f[i])
    flag = torch.tensor([int(ok and ok_g and ok_1)], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("fused==nccl %s, graph replay==eager %s, ==tp1 %s" % (ok, ok_g, ok_1))
        print("TP_NVLINK_OK" if int(flag[0]) == 1 else "TP_NVLINK_MISMATCH")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if int(flag[0]) == 1 else 1)


if __name__ == "__main__":
    main()
