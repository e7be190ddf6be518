"""Multi-GPU correctness checks of the symmetric-memory kernels (csrc/ws_collective.cu), run under torchrun by
tests/test_gpu_multi.py when the box has >= 2 (>= 4) GPUs:

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/multi_rank_checks.py

* ws_allreduce_sum_f32, both kernels (peer loads/stores; multimem where the switch offers multicast), at sizes that
  are NOT a multiple of 4 x world (the slice tail of `per = ceil(n4 / world)`), incl. sizes smaller than the world,
  against torch.distributed.all_reduce: max relative error, identical bits on every rank, elements past `numel`
  untouched;
* dist.ShardedAdam (reduce-scatter + Adam + all-gather in one kernel) against all-reduce + fused Adam: parameters bit
  for bit on every rank;
* the vote exchange of config 3 (dist.make_bucket over used_count) and dist.gather_frames.
Prints one "OK <name>" line per check on rank 0; exits non-zero on the first failure.
"""
import os
import sys

import torch
import torch.distributed as tdist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import dist as wdist, optim as woptim  # noqa: E402

rank, world, local = wdist.init_from_env("nccl")
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
FAIL = []


def agree(flag):
    t = torch.tensor([1.0 if flag else 0.0], device=dev)
    tdist.all_reduce(t, op=tdist.ReduceOp.MIN)
    return bool(t.item())


def report(name, ok, detail=""):
    ok = agree(ok)
    if rank == 0:
        print("%s %s %s" % ("OK" if ok else "FAIL", name, detail), flush=True)
    if not ok:
        FAIL.append(name)


def check_allreduce(kernel, sizes):
    os.environ["WS_ALLREDUCE_KERNEL"] = kernel
    cap = max(sizes) + 64
    sb = wdist.SymmetricBucket(cap, dev)
    if kernel == "multimem" and not sb.multicast:
        report("allreduce[%s]" % kernel, True, "(no multicast on this box: skipped)")
        return
    for n in sizes:
        g = torch.Generator(device=dev).manual_seed(1000 * n + rank)
        src = torch.randn(sb.numel, generator=g, device=dev)
        src[::7] *= 1e5
        sb.tensor.copy_(src)
        ref = src.clone()
        tdist.barrier()
        sb.all_reduce_(n)
        n4 = (n + 3) // 4 * 4
        tdist.all_reduce(ref[:n4], op=tdist.ReduceOp.SUM)
        torch.cuda.synchronize()
        scale = float(ref[:n4].abs().max()) if n4 else 1.0
        err = float((sb.tensor[:n4] - ref[:n4]).abs().max()) / scale if n4 else 0.0
        tail_ok = torch.equal(sb.tensor[n4:], src[n4:])
        words = sb.tensor[:n4].view(torch.int32).to(torch.int64)
        chk = torch.stack([words.sum(), (words * (torch.arange(n4, device=dev) % 8191 + 1)).sum()]) if n4 else \
            torch.zeros(2, dtype=torch.int64, device=dev)
        allc = [torch.zeros_like(chk) for _ in range(world)]
        tdist.all_gather(allc, chk)
        same = all(torch.equal(c, allc[0]) for c in allc)
        report("allreduce[%s] n=%d" % (kernel, n), err < 1e-5 and tail_ok and same,
               "rel err %.2e, tail untouched %s, identical across ranks %s" % (err, tail_ok, same))
    del sb


def check_sharded_adam(P):
    M = 16
    shapes = [("xyz", (P, 3), 0.00016, 0), ("f_dc", (P, 1, 3), 0.0025, 6 * P), ("f_rest", (P, 15, 3), 0.000125, 9 * P),
              ("opacity", (P, 1), 0.05, 54 * P), ("scaling", (P, 3), 0.005, 55 * P), ("rotation", (P, 4), 0.001, 58 * P)]
    gen = torch.Generator(device=dev).manual_seed(1)
    init = {n: torch.randn(s, device=dev, generator=gen) for n, s, _, _ in shapes}
    bucket = wdist.SymmetricBucket(wdist.full_bucket_numel(P, M), dev)
    n_train = wdist.train_bucket_numel(P, M)
    sh = wdist.ShardedAdam([{"name": n, "tensor": init[n], "lr": lr, "grad_offset": off} for n, s, lr, off in shapes],
                           bucket, eps=1e-15)
    ref = {n: torch.nn.Parameter(init[n].clone()) for n, _, _, _ in shapes}
    opt = woptim.Adam([{"params": [ref[n]], "lr": lr, "name": n} for n, _, lr, _ in shapes], lr=0.0, eps=1e-15)
    ggen = torch.Generator(device=dev).manual_seed(100 + rank)
    ok = True
    for _ in range(3):
        lg = torch.randn(n_train, device=dev, generator=ggen) * 1e-3
        bucket.tensor[:n_train].copy_(lg)
        sh.step()
        bucket.tensor[:n_train].copy_(lg)
        bucket.all_reduce_(n_train)
        for n, s, _, off in shapes:
            ref[n].grad = bucket.tensor[off:off + ref[n].numel()].view(s)
        opt.step()
        torch.cuda.synchronize()
        ok = ok and all(torch.equal(sh.params[n].detach(), ref[n].detach()) for n, _, _, _ in shapes)
    report("sharded_adam P=%d (%s)" % (P, "multimem" if sh.param_mc else "peers"), ok, "bit-identical to all-reduce + Adam")


def check_votes_and_frames():
    rows, P = 13, 4000
    votes, reduce_votes, how = wdist.make_bucket(rows * P, dev)
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    votes[:rows * P].copy_(torch.rand(rows * P, generator=g, device=dev))
    ref = votes[:rows * P].clone()
    tdist.barrier()
    reduce_votes(rows * P)
    tdist.all_reduce(ref, op=tdist.ReduceOp.SUM)
    torch.cuda.synchronize()
    report("vote exchange (%s)" % how, float((votes[:rows * P] - ref).abs().max()) < 1e-5)
    frames = [torch.full((3, 4, 5), float(k), device=dev) for k in range(rank, 11, world)]
    got = wdist.gather_frames(frames, 11, rank, world)
    ok = True
    if rank == 0:
        ok = len(got) == 11 and all(float(f.mean()) == float(k) for k, f in enumerate(got))
    report("gather_frames", ok)


sizes = [4, 8, 4 * world - 4, 4 * world + 4, 4 * (world * 1000 + 1), 4 * (world * 1000 + world - 1), 1000003 * 4,
         6200004, 4 * 5 + 3]   # the last one is not a multiple of 4: rounded up by all_reduce_
check_allreduce("peers", sizes)
check_allreduce("multimem", sizes)
os.environ.pop("WS_ALLREDUCE_KERNEL", None)
check_sharded_adam(20000)
check_votes_and_frames()
tdist.barrier()
tdist.destroy_process_group()
sys.exit(1 if FAIL else 0)
