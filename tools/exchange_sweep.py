"""Exchange kernel sweep (N GPUs): ws_allreduce_sum_f32 over the training gradient bucket of P Gaussians for the grid /
unroll / threads settings of `ws_allreduce_tune`, both kernels, against NCCL all_reduce; CUDA events, max over ranks,
barriers included (what a training step pays).  Link traffic per direction for S bytes on W GPUs: multimem S (each
GPU's copy leaves once, the reduced slices come back once), peers 2 (W-1)/W S.

    torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/exchange_sweep.py [P ...]   -> JSON lines (rank 0)
"""
import json
import os
import sys

import torch
import torch.distributed as tdist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import dist as wdist, lib as wlib  # noqa: E402

rank, world, local = wdist.init_from_env("nccl")
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
L = wlib.load()


def timed(fn, n=20):
    for _ in range(3):
        fn()
    tdist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return wdist.max_over_ranks(e0.elapsed_time(e1) / n, dev)


for P in [int(a) for a in sys.argv[1:]] or [1000000]:
    n = wdist.train_bucket_numel(P, 16)
    S = n * 4
    x = torch.randn(n, device=dev)
    res = {"P": P, "world": world, "bytes": S, "nccl_ms": timed(lambda: tdist.all_reduce(x))}
    for kernel in ("multimem", "peers"):
        os.environ["WS_ALLREDUCE_KERNEL"] = kernel
        sb = wdist.SymmetricBucket(n, dev)
        if kernel == "multimem" and not sb.multicast:
            continue
        sb.tensor.normal_()
        for ctas in (74, 148, 296, 592):
            for unroll, threads in ((4, 512), (8, 512), (8, 256), (2, 512)) if kernel == "multimem" else ((4, 512),):
                wlib.check(L.ws_allreduce_tune(unroll, threads), "ws_allreduce_tune")
                sb.num_ctas = ctas
                ms = timed(lambda: sb.all_reduce_(n))
                traffic = S if kernel == "multimem" else 2 * (world - 1) / world * S
                res["%s ctas=%d unroll=%d threads=%d" % (kernel, ctas, unroll, threads)] = \
                    {"ms": ms, "link_gbs_per_direction": traffic / ms / 1e6}
        del sb
    L.ws_allreduce_tune(4, 512)
    if rank == 0:
        print(json.dumps(res), flush=True)
tdist.destroy_process_group()
