"""CPU, world_size 2 over gloo: the host logic of the N>1 path (view sharding, ONE all_reduce of the
gradient bucket, vote accumulators, densification statistics, frame gather, max-over-ranks)."""
import os
import socket

import torch
import torch.distributed as tdist
import torch.multiprocessing as mp

from wheat_3dgs_b200 import dist as wdist


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    r, w, _ = wdist.init_from_env("gloo")
    assert (r, w) == (rank, world) and wdist.world_size() == world
    P, M = 50, 16
    views = wdist.shard(range(7), rank, world)
    # every "view" contributes a known gradient: view id + 1 in the training set, rank marker in the tail
    bucket = torch.zeros(wdist.full_bucket_numel(P, M))
    for v in views:
        bucket[:wdist.train_bucket_numel(P, M)] += float(v + 1)
    bucket[wdist.train_bucket_numel(P, M):] = float(100 + rank)
    wdist.allreduce_train_grads(bucket, P, M)
    s = wdist.split_bucket(bucket, P, M)
    ok = bool((s["dL_dsh"] == 28.0).all() and (s["dL_dmeans3D"] == 28.0).all() and (s["dL_drotations"] == 28.0).all())
    ok &= bool((s["dL_dcov3D"] == 100 + rank).all())  # not part of the collective
    votes = torch.full((3, P), float(rank + 1))
    wdist.allreduce_votes(votes)
    ok &= bool((votes == 3.0).all())
    acc, den, rad = torch.full((P, 1), 1.0 + rank), torch.ones(P, 1), torch.full((P,), 5.0 * (rank + 1))
    wdist.allreduce_densify_stats(acc, den, rad)
    ok &= bool((acc == 3.0).all() and (den == 2.0).all() and (rad == 10.0).all())
    frames = wdist.gather_frames([torch.full((2, 2), float(k)) for k in wdist.shard(range(5), rank, world)], 5, rank, world)
    if rank == 0:
        ok &= [int(f[0, 0]) for f in frames] == [0, 1, 2, 3, 4]
    else:
        ok &= frames is None
    ok &= wdist.max_over_ranks(10.0 + rank, torch.device("cpu")) == 11.0
    q.put((rank, ok))
    tdist.destroy_process_group()


def test_two_ranks_over_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=180) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
