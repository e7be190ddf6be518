"""N-GPU check of the in-place symmetric-memory all-reduce (csrc/ws_collective.cu) against NCCL, with timing.
torchrun --nproc-per-node N tools/allreduce_check.py"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import dist as wdist  # noqa: E402

rank, world, local = wdist.init_from_env("nccl")
dev = torch.device("cuda", local)
n = 62 * 1000000
gen = torch.Generator(device=dev).manual_seed(100 + rank)
x = torch.randn(n, device=dev, generator=gen)


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return wdist.max_over_ranks(e0.elapsed_time(e1) / iters, dev)


ref = x.clone()
dist.all_reduce(ref)
for mode in ("multimem", "peers"):
    if mode == "peers":
        os.environ["WS_ALLREDUCE_PEERS"] = "1"
    sb = wdist.SymmetricBucket(n, dev)
    if mode == "multimem" and not sb.multicast:
        continue
    sb.tensor.copy_(x)
    sb.all_reduce_()
    torch.cuda.synchronize()
    err = (sb.tensor - ref).abs().max().item()
    rel = err / ref.abs().max().item()
    # all ranks must hold the same bits
    mine = sb.tensor.view(torch.int32).sum(dtype=torch.int64)
    lo, hi = mine.clone(), mine.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    for ctas in (0, 64, 296, 592):
        sb.num_ctas = ctas
        ms = timeit(sb.all_reduce_)
        if rank == 0:
            print("%s ctas=%d: %.3f ms (%.0f GB/s algorithmic)  max err vs NCCL %.3g (rel %.2g)  identical on all ranks: %s"
                  % (mode, ctas or 148, ms, n * 4 / ms / 1e6, err, rel, bool(lo == hi)), flush=True)
ms = timeit(lambda: dist.all_reduce(x))
if rank == 0:
    print("NCCL all_reduce: %.3f ms (%.0f GB/s algorithmic)" % (ms, n * 4 / ms / 1e6), flush=True)
dist.destroy_process_group()
