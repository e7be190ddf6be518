"""View / frame / mask sharding across the GPUs of one box (one process per GPU).

The reference is single-GPU (utils/general_utils.py:133 pins cuda:0) and does one view per
iteration (train_vanilla_3dgs.py:65).  The hot path shards by independent units with ONE
exchange step:

  training views   rank r takes views r, r+N, ... of the epoch; every rank holds all Gaussians;
                   the per-Gaussian gradients of the N views of a step are summed with one NCCL
                   all_reduce over the flat bucket ops._backward() writes
                   ([means3D 3 | means2D 3 | sh 3M | opacity 1 | scales 3 | rotations 4] x P floats,
                   the training-time gradient set, contiguous).
  label votes      (view, mask) units round-robin; used_count[(num_obj+1) x P] accumulated locally
                   across the rank's views in one persistent buffer, all-reduced ONCE at the end
                   (render_360.py:62-69 / run_3d_seg.py:97 add per-view counts on the host).
  orbit frames     frames r::N, no exchange (render_360, utils/wheatgs_helper.py:282).

Works with the `nccl` backend on CUDA tensors and with `gloo` on CPU tensors (the CPU tests run
the host logic at world_size 2 under gloo).
"""
import os

import torch
import torch.distributed as dist

TRAIN_GRAD_FLOATS = lambda M: 3 + 3 + 3 * M + 1 + 3 + 4  # noqa: E731  per Gaussian
DEFAULT_SPLIT = ""   # "permille,peer_ctas" for the hybrid all-reduce (multimem + peer path in one launch); "" = switch only


def init_from_env(backend=None):
    """Initialises torch.distributed from RANK / WORLD_SIZE / MASTER_* (torchrun).  Returns
    (rank, world_size, local_rank).  Single-process runs return (0, 1, 0) without a group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def world_size():
    return dist.get_world_size() if dist.is_initialized() else 1


def shard(units, rank, world):
    """Round-robin partition of independent units (views, frames, (view, mask) pairs)."""
    return list(units)[rank::world]


def balanced_schedule(costs, world, n_steps):
    """Lock-step data parallelism ends every step with the exchange, so a step lasts as long as its slowest view.
    Views differ in cost (the instance count R varies by +-25 % over the overhead views), hence views of SIMILAR
    cost are put into the same step: the views are sorted by cost and dealt out `world` at a time, cycling through
    the sorted list (every view is visited equally often; a trainer shuffles the ORDER of the steps, not their
    composition, and refreshes the costs from the previous epoch's R).  Returns schedule[step][rank] = view index."""
    order = sorted(range(len(costs)), key=lambda v: (-float(costs[v]), v))
    n = len(order)
    return [[order[(s * world + r) % n] for r in range(world)] for s in range(n_steps)]


def shard_counts(n_units, world):
    """Units per rank of shard(range(n_units), r, world) for every r."""
    return [len(range(r, n_units, world)) for r in range(world)]


def train_bucket_numel(P, M):
    return TRAIN_GRAD_FLOATS(M) * P


def full_bucket_numel(P, M):
    """ops._backward() layout: the training set followed by [colors 3 | cov3D 6]."""
    return (TRAIN_GRAD_FLOATS(M) + 9) * P


def split_bucket(bucket, P, M):
    """Views into the flat gradient bucket, keyed like the reference's returned gradients."""
    sizes = (("dL_dmeans3D", 3), ("dL_dmeans2D", 3), ("dL_dsh", 3 * M), ("dL_dopacity", 1), ("dL_dscales", 3),
             ("dL_drotations", 4), ("dL_dcolors", 3), ("dL_dcov3D", 6))
    out, off = {}, 0
    for name, s in sizes:
        out[name] = bucket[off:off + s * P]
        off += s * P
    out["dL_dsh"] = out["dL_dsh"].view(P, M, 3)
    for name, s in sizes:
        if name != "dL_dsh":
            out[name] = out[name].view(P, s)
    return out


def allreduce_train_grads(bucket, P, M, async_op=False):
    """SUM over ranks of the training gradient set (62*P floats at SH degree 3) in ONE collective."""
    if world_size() == 1:
        return None
    return dist.all_reduce(bucket[:train_bucket_numel(P, M)], op=dist.ReduceOp.SUM, async_op=async_op)


class SymmetricBucket:
    """The gradient bucket in symmetric memory (every rank maps every peer's copy; with NVSwitch also one
    multicast address for all of them), summed over ranks in place by ONE kernel of libwheatsplat
    (csrc/ws_collective.cu: multimem.ld_reduce / multimem.st, or peer loads and stores) on the CURRENT stream,
    bracketed by the symmetric-memory barrier -- no hand-off to NCCL's stream.  NCCL stays the plumbing
    (rendezvous, everything else in this module) and the fallback of `make_bucket` when symmetric memory
    cannot be set up.  `tensor` is what ops._backward() writes into."""

    def __init__(self, numel, device, group=None):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm
        from . import lib as _lib
        self._lib, self._C = _lib, C
        group = group if group is not None else dist.group.WORLD
        self.numel = (int(numel) + 3) // 4 * 4
        self.tensor = symm.empty(self.numel, dtype=torch.float32, device=device)
        self.hdl = symm.rendezvous(self.tensor, group.group_name)
        self.rank, self.world = self.hdl.rank, self.hdl.world_size
        # Link traffic per direction for a bucket of S bytes on W GPUs: multicast (1 + 1/W) S, peer loads/stores
        # 2 (W-1)/W S -> multicast from W = 4 up (measured at W = 2: peers 0.39 ms, multicast 0.62 ms, NCCL 0.49 ms
        # for 248 MB).  WS_ALLREDUCE_KERNEL=peers|multimem overrides.
        self.multicast = int(self.hdl.multicast_ptr or 0)   # 0: no NVSwitch multicast
        want = os.environ.get("WS_ALLREDUCE_KERNEL", "multimem" if self.world >= 4 else "peers")
        if want == "peers" or os.environ.get("WS_ALLREDUCE_PEERS"):
            self.multicast = 0
        self.peers = (C.c_void_p * self.world)(*[int(p) for p in self.hdl.buffer_ptrs])
        self.num_ctas = int(os.environ.get("WS_ALLREDUCE_CTAS", "0"))
        # share of every slice that goes through the switch, the rest by peer loads / stores at the same time
        # (ws_allreduce_split; "permille,peer_ctas")
        split = os.environ.get("WS_ALLREDUCE_SPLIT", DEFAULT_SPLIT)
        if self.multicast and split:
            pm, pc = (int(x) for x in split.split(","))
            _lib.check(_lib.load().ws_allreduce_split(pm, pc), "ws_allreduce_split")

    def all_reduce_(self, numel=None):
        """SUM over ranks of tensor[:numel] (default: all of it), in place, on the current stream."""
        n = self.numel if numel is None else (int(numel) + 3) // 4 * 4
        L = self._lib.load()
        self.hdl.barrier(channel=0)
        with torch.cuda.device(self.tensor.device):
            st = self._C.c_void_p(torch.cuda.current_stream().cuda_stream)
            self._lib.check(L.ws_allreduce_sum_f32(self.rank, self.world, n, self._C.c_void_p(self.multicast or None),
                                                   self.peers, self.num_ctas, st), "ws_allreduce_sum_f32")
        self.hdl.barrier(channel=1)
        return self.tensor


def make_bucket(numel, device):
    """(tensor, reduce) for the gradient bucket: a SymmetricBucket summed by libwheatsplat's own kernel when the
    group spans GPUs with peer access, else a plain tensor summed by one NCCL all_reduce.  WS_ALLREDUCE=nccl
    forces the latter."""
    if world_size() > 1 and torch.device(device).type == "cuda" and os.environ.get("WS_ALLREDUCE", "") != "nccl":
        try:
            sb = SymmetricBucket(numel, device)
            return sb.tensor, sb.all_reduce_, "libwheatsplat multimem kernel" if sb.multicast else "libwheatsplat peer kernel"
        except Exception as e:  # noqa: BLE001 - no peer access / symmetric memory on this box
            if os.environ.get("WS_ALLREDUCE", "") == "symm":
                raise
            print("wheat_3dgs_b200.dist: symmetric memory unavailable (%s); NCCL all_reduce" % str(e)[:200], flush=True)
    t = torch.empty((numel,), dtype=torch.float32, device=device)

    def reduce(n=None):
        if world_size() > 1:
            dist.all_reduce(t if n is None else t[:n], op=dist.ReduceOp.SUM)
        return t
    return t, reduce, "NCCL all_reduce"


def allreduce_votes(used_count, async_op=False):
    """SUM over ranks of the label-vote accumulator used_count[(num_obj+1), P]."""
    if world_size() == 1:
        return None
    return dist.all_reduce(used_count, op=dist.ReduceOp.SUM, async_op=async_op)


def allreduce_densify_stats(xyz_gradient_accum, denom, max_radii2D):
    """Densification statistics (scene/gaussian_model.py:461-463): SUM, SUM, MAX, IN PLACE on every rank.
    The statistics are cumulative over a densification interval, so call this EXACTLY ONCE per interval, immediately
    before densify_and_prune reads them (which then resets them): a second call would add the already summed values
    again (x world_size)."""
    if world_size() == 1:
        return
    dist.all_reduce(xyz_gradient_accum, op=dist.ReduceOp.SUM)
    dist.all_reduce(denom, op=dist.ReduceOp.SUM)
    dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX)


def gather_frames(frame, n_frames, rank, world):
    """Frames rendered as shard(range(n_frames)) gathered in frame order on rank 0 (None elsewhere)."""
    if world == 1:
        return [frame] if not isinstance(frame, list) else frame
    mine = frame if isinstance(frame, list) else [frame]
    out = [None] * world
    dist.all_gather_object(out, [f.cpu() for f in mine])
    if rank != 0:
        return None
    frames = [None] * n_frames
    for r in range(world):
        for k, f in zip(range(r, n_frames, world), out[r]):
            frames[k] = f
    return frames


def max_over_ranks(value, device):
    """Max over ranks of a python float (timings are reported as the slowest rank's)."""
    if world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class ShardedAdam:
    """Exchange + optimiser step of data-parallel training as ONE kernel (csrc/ws_collective.cu `sharded_adam`):
    the replicated parameters live in one symmetric-memory buffer, rank r reduces the r-th slice of the gradient
    bucket through the switch, applies torch.optim.Adam's step to it with its slice of the moments and stores the
    new parameters into every replica.  Same numbers as `SymmetricBucket.all_reduce_()` followed by
    `optim.Adam.step()` on every rank, bit for bit; 1/W of the Adam traffic per rank and no gradient broadcast.

    groups: [{"name", "tensor" (initial value, any device), "lr", "grad_offset" (floats into the bucket)}], in the
    order they are laid out; every numel and offset a multiple of 4.  `params[name]` are nn.Parameters viewing the
    symmetric buffer (what the model trains); `bucket` is the SymmetricBucket the backward writes into.

    The flat buffers have a fixed size: after densify_and_prune (a new P) build a NEW ShardedAdam (and bucket) from
    the new tensors with `rebuild()`; P must be a multiple of 4 (pad the model with culled rows otherwise)."""

    def __init__(self, groups, bucket, betas=(0.9, 0.999), eps=1e-8):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm
        from . import lib as _lib
        from .lib import WsShardSegment
        if not isinstance(bucket, SymmetricBucket):
            raise RuntimeError("ShardedAdam needs the gradient bucket in symmetric memory (dist.SymmetricBucket)")
        self._lib, self._C, self._Seg = _lib, C, WsShardSegment
        self.bucket, self.betas, self.eps = bucket, betas, float(eps)
        dev = bucket.tensor.device
        self.numel = sum(int(g["tensor"].numel()) for g in groups)
        if any(int(g["tensor"].numel()) % 4 or int(g["grad_offset"]) % 4 for g in groups):
            raise RuntimeError("ShardedAdam: group sizes and gradient offsets must be multiples of 4 floats")
        self.flat = symm.empty(self.numel, dtype=torch.float32, device=dev)
        self.hdl = symm.rendezvous(self.flat, dist.group.WORLD.group_name)
        self.params, self.groups, off = {}, [], 0
        for g in groups:
            n = int(g["tensor"].numel())
            view = self.flat[off:off + n].view(g["tensor"].shape)
            view.copy_(g["tensor"].detach().to(dev, torch.float32))
            self.params[g["name"]] = torch.nn.Parameter(view.requires_grad_(True))
            self.groups.append({"name": g["name"], "begin": off, "end": off + n, "grad_offset": int(g["grad_offset"]),
                                "lr": float(g["lr"])})
            off += n
        self.exp_avg = torch.zeros(self.numel, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(self.numel, dtype=torch.float32, device=dev)
        self.step_count = 0
        use_mc = bool(bucket.multicast) and bool(self.hdl.multicast_ptr)
        self.grad_mc = bucket.multicast if use_mc else 0
        self.param_mc = int(self.hdl.multicast_ptr or 0) if use_mc else 0
        self.param_peers = (C.c_void_p * bucket.world)(*[int(p) for p in self.hdl.buffer_ptrs])
        self.num_ctas = int(os.environ.get("WS_SHARDED_ADAM_CTAS", "0"))
        self.hdl.barrier(channel=0)   # every replica initialised before anyone steps

    def rebuild(self, groups, bucket, moments=None):
        """A new ShardedAdam for the post-densification tensors; `moments` = {name: (exp_avg, exp_avg_sq)} rows to
        carry over (the reference keeps the moments of surviving rows, scene/gaussian_model.py:296-345)."""
        new = ShardedAdam(groups, bucket, betas=self.betas, eps=self.eps)
        new.step_count = self.step_count
        if moments:
            for g in new.groups:
                if g["name"] in moments:
                    m, v = moments[g["name"]]
                    new.exp_avg[g["begin"]:g["end"]].copy_(m.reshape(-1))
                    new.exp_avg_sq[g["begin"]:g["end"]].copy_(v.reshape(-1))
        return new

    def set_lr(self, name, lr):
        for g in self.groups:
            if g["name"] == name:
                g["lr"] = float(lr)

    @torch.no_grad()
    def step(self):
        """After every rank's backward has written its gradients into the bucket (same stream)."""
        C, b = self._C, self.bucket
        self.step_count += 1
        segs = (self._Seg * len(self.groups))(*[self._Seg(g["begin"], g["end"], g["grad_offset"], g["lr"])
                                                for g in self.groups])
        L = self._lib.load()
        b.hdl.barrier(channel=0)      # all gradients written, all reads of the old parameters done
        with torch.cuda.device(self.flat.device):
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            self._lib.check(L.ws_sharded_adam_step(b.rank, b.world, self.numel, len(self.groups), segs,
                                                   C.c_void_p(self.grad_mc or None), b.peers,
                                                   C.c_void_p(self.param_mc or None), self.param_peers,
                                                   self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.step_count,
                                                   float(self.betas[0]), float(self.betas[1]), self.eps, self.num_ctas, st),
                            "ws_sharded_adam_step")
        b.hdl.barrier(channel=1)      # every replica holds the new parameters, the bucket may be rewritten
