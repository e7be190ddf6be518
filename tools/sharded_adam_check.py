"""Fused exchange + Adam step (dist.ShardedAdam, csrc/ws_collective.cu `sharded_adam`) against the two-kernel
form (SymmetricBucket.all_reduce_ + optim.Adam.step on every rank): parameters must agree bit for bit on every
rank after every step; then both are timed with CUDA events (max over ranks).

    torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/sharded_adam_check.py [P]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import dist as wdist, optim as woptim  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
M = 16
rank, world, local = wdist.init_from_env()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
# parameter space order and the gradient bucket's layout (ops._backward: means3D 3 | means2D 3 | sh: dc 3, rest 45 | ...)
SHAPES = [("xyz", (P, 3), 0.00016, 0), ("f_dc", (P, 1, 3), 0.0025, 6 * P), ("f_rest", (P, 15, 3), 0.000125, 9 * P),
          ("opacity", (P, 1), 0.05, 54 * P), ("scaling", (P, 3), 0.005, 55 * P), ("rotation", (P, 4), 0.001, 58 * P)]
gen = torch.Generator(device=dev).manual_seed(1)
init = {n: torch.randn(s, device=dev, generator=gen) for n, s, _, _ in SHAPES}
bucket = wdist.SymmetricBucket(wdist.full_bucket_numel(P, M), dev)
n_train = wdist.train_bucket_numel(P, M)
sh = wdist.ShardedAdam([{"name": n, "tensor": init[n], "lr": lr, "grad_offset": off} for n, s, lr, off in SHAPES], bucket,
                       eps=1e-15)
ref = {n: torch.nn.Parameter(init[n].clone()) for n, _, _, _ in SHAPES}
opt = woptim.Adam([{"params": [ref[n]], "lr": lr, "name": n} for n, _, lr, _ in SHAPES], lr=0.0, eps=1e-15)
ggen = torch.Generator(device=dev).manual_seed(100 + rank)
ok = True
for step in range(3):
    local_grads = torch.randn(n_train, device=dev, generator=ggen) * 1e-3
    bucket.tensor[:n_train].copy_(local_grads)
    sh.step()
    bucket.tensor[:n_train].copy_(local_grads)
    bucket.all_reduce_(n_train)
    for n, s, _, off in SHAPES:
        numel = ref[n].numel()
        ref[n].grad = bucket.tensor[off:off + numel].view(s)
    opt.step()
    torch.cuda.synchronize()
    for n, _, _, _ in SHAPES:
        same = torch.equal(sh.params[n].detach(), ref[n].detach())
        ok = ok and same
        if not same:
            d = (sh.params[n].detach() - ref[n].detach()).abs().max().item()
            print("rank %d step %d %s differs, max abs %g" % (rank, step, n, d), flush=True)
flag = torch.tensor([1.0 if ok else 0.0], device=dev)
torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN)


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.distributed.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return wdist.max_over_ranks(e0.elapsed_time(e1) / n, dev)


def two_kernels():
    bucket.all_reduce_(n_train)
    opt.step()


t_fused = timed(sh.step)
t_two = timed(two_kernels)
sweep = {}
for ctas in (74, 296, 592):
    sh.num_ctas = ctas
    sweep[ctas] = timed(sh.step)
sh.num_ctas = 0
if rank == 0:
    print("%d GPUs, P=%d (%s): fused reduce+Adam+broadcast %.3f ms; all-reduce + Adam %.3f ms; bit-identical on all ranks: %s"
          % (world, P, "multimem" if sh.param_mc else "peers", t_fused, t_two, bool(flag.item())), flush=True)
    print("   CTAs (default 148): " + "  ".join("%d: %.3f ms" % kv for kv in sweep.items()), flush=True)
torch.distributed.destroy_process_group()
sys.exit(0 if flag.item() else 1)
