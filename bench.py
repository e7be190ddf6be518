#!/usr/bin/env python
"""bench.py -- the splat hot path on BASELINE.json's config 2.

Workload (`config.workload` = "cfg2"): synthetic plot_461-shaped scene, 1 M Gaussians, SH degree 3,
36 overhead views at 2048x1536 (wheat_3dgs_b200/scene.py, seeds 0/1/2).  One STEP = the rasterizer
part of one training iteration on one view: forward (preprocess, key emission, radix sort, tile
ranges, blend with depth+alpha) + backward (blend backward, preprocess backward) -- exactly what
train_vanilla_3dgs.py:73,80 spends in the three submodules.  With N GPUs every rank takes its own
views (weak scaling: N views per step, Gaussians replicated) and the step ends with ONE NCCL
all_reduce of the 62*P-float gradient bucket.

  value      views/s over all ranks, inputs resident in HBM, C-ABI operator calls
             (ops.rasterize_gaussians + ops.rasterize_gaussians_backward)
  e2e        views/s through the drop-in GaussianRasterizer + autograd, with this step's ground-truth
             image (3xHxW fp32) and camera copied from pinned host memory and the loss read back
             to the host inside the timed region (each step's 4 bytes are copied behind the step and
             read one step later, the last before the closing event: ResultReader, both arms)
  roofline   the dominant stage, timed with CUDA events recorded by the library on the launching
             stream in a second pass over the same steps
  cpu_baseline  the CPU oracle (oracle/splat_oracle.c, a port) on a bounded sample, rank 0, N=1

`--impl reference` runs the UNMODIFIED reference CUDA rasterizer (oracle/_ref/libref_dgr.so, built
from /root/reference by oracle/build_ref.py) on the same GPU, same views, same timing: the reference
has no CPU implementation of this path (it is CUDA-only), so the reference arm is its stock CUDA
path -- the number the north star's ">= 2x" is against.  `--impl reference-cpu` times the CPU
oracle port instead.  Neither touches libwheatsplat.so.
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

WORKLOADS = {
    # name: (P, W, H, n_views)
    "cfg2": (1000000, 2048, 1536, 36),
    "cfg1": (50000, 512, 512, 1),
    "tiny": (20000, 320, 240, 4),  # plumbing check only (tests)
}
SH_DEGREE = 3
BG = np.array([0.0, 0.0, 0.0], np.float32)  # train_vanilla_3dgs.py:50 (black unless --white_background)


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        if os.environ.get("WS_NO_CLOCKS"):
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-i", str(self.index),
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def wait_started(self, n=2, timeout=5.0):
        """nvidia-smi takes a while to attach and holds up CUDA API calls while it does: the timed region
        must not start before it is polling steadily."""
        t0 = time.time()
        while self.proc is not None and len(self.lines) < n and time.time() - t0 < timeout:
            time.sleep(0.02)

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for ts, ln in self.lines:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(self.NAMES, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": sorted(reasons), "samples": len(sm)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": float(d["hbm_gbs"]), "sm_max_mhz": float(d.get("sm_max_mhz", 1965.0)), "source": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}


# ----------------------------------------------------------------------------- workload
def make_workload(name):
    from wheat_3dgs_b200 import scene
    P, W, H, nv = WORKLOADS[name]
    sc = scene.synthetic_scene(P, seed=0, sh_degree=SH_DEGREE)
    if nv >= 36:
        cams = scene.overhead_cameras(W, H, seed=1)
    else:
        cams = scene.overhead_cameras(W, H, seed=1)[17:17 + nv]
    return sc, cams, (P, W, H, len(cams))


def gt_images(n, H, W, seed=2):
    """Pinned host 'ground-truth' images (seeded noise; the loss value is irrelevant to timing)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    imgs = []
    for _ in range(n):
        t = torch.rand((3, H, W), generator=g, dtype=torch.float32)
        imgs.append(t.pin_memory() if torch.cuda.is_available() else t)
    return imgs


class InputPrefetcher:
    """The harness's input pipeline, shared by both arms (a data loader's prefetch).  Step i's camera (35
    floats) + ground-truth image (3xHxW fp32) travel from pinned host memory on a copy stream while step
    i-1 computes (two device slots); the compute stream waits for the copy's event.  Every step's H2D
    copy is issued and completes inside the timed region."""

    def __init__(self, dev, H, W, cam_host, gts, view_of):
        import torch
        self.torch, self.cam_host, self.gts, self.view_of = torch, cam_host, gts, view_of
        self.stream = torch.cuda.Stream(device=dev)
        self.slots = [{"cam": torch.empty(35, dtype=torch.float32, device=dev),
                       "gt": torch.empty((3, H, W), dtype=torch.float32, device=dev),
                       "ready": torch.cuda.Event(), "free": torch.cuda.Event()} for _ in range(2)]
        for sl in self.slots:
            sl["free"].record()

    def prefetch(self, i):
        v = self.view_of(i)
        sl = self.slots[i & 1]
        with self.torch.cuda.stream(self.stream):
            self.stream.wait_event(sl["free"])  # the step that last read this slot has finished with it
            sl["cam"].copy_(self.cam_host[v], non_blocking=True)
            sl["gt"].copy_(self.gts[v % len(self.gts)], non_blocking=True)
            sl["ready"].record(self.stream)

    def take(self, i, last):
        """Inputs of step i (waits for their copy); starts the copy of step i+1's unless `last`."""
        sl = self.slots[i & 1]
        self.torch.cuda.current_stream().wait_event(sl["ready"])
        if not last:
            self.prefetch(i + 1)
        return sl

    @staticmethod
    def release(sl):
        sl["free"].record()


class ResultReader:
    """The harness's result pipeline, shared by both arms.  Every step's loss (4 bytes) is copied to pinned host
    memory behind the step's kernels and READ on the host one step later, once the next step has been queued
    (a training loop's logging, one step delayed), so the device never drains while the host formats a number.
    Every step's value is read inside the timed region: `flush()` reads the last one before the closing event."""

    def __init__(self, torch):
        self.torch = torch
        self.host = torch.empty(2, dtype=torch.float32).pin_memory()
        self.ev = [torch.cuda.Event(), torch.cuda.Event()]
        self.pending = None
        self.n = 0
        self.values = []

    def push(self, loss):
        k = self.n & 1
        self.host[k:k + 1].copy_(loss.detach().reshape(1), non_blocking=True)
        self.ev[k].record()
        prev, self.pending = self.pending, k
        self.n += 1
        if prev is not None:
            self._read(prev)

    def _read(self, k):
        self.ev[k].synchronize()
        self.values.append(float(self.host[k]))

    def flush(self):
        if self.pending is not None:
            self._read(self.pending)
            self.pending = None


def stage_model(P, V, R, T, HW, M, npass, n_eval, n_contrib):
    """Algorithmic HBM bytes / flops per view of every stage (DESIGN.md section 4 derives them)."""
    c = 12 * M  # SH bytes per Gaussian
    by = {
        # depth-first binning (default): preprocess also writes the (depth bits, id) pair; "sort" =
        # the depth sort of P 8-byte pairs (4 passes) + the tile sort of R 8-byte pairs (npass passes)
        "preprocess": P * (44 + 8 + 8) + V * (c + 49),
        "scan": 8 * ((P + 255) // 256),
        "emit_keys": 12 * P + 8 * ((P + 255) // 256) + 24 * V + 8 * R,
        "sort": (4 * P + 4 * 16 * P) + (4 * R + npass * 16 * R),
        "ranges_pack": 8 * R + 48 * R + 8 * T,
        "render_fwd": 48 * R + 24 * HW,
        "render_bwd": 48 * R + 32 * HW + 40 * V,
        "preprocess_bwd": P * (4 + 48) + V * (44 + c) + P * (c + 92),
    }
    fl = {"render_fwd": 11 * n_eval + 17 * n_contrib, "render_bwd": 13 * n_eval + 60 * n_contrib}
    return by, fl


def to_dev(sc, cams, dev):
    import torch
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    g = {k: t(sc[k]) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    cd = []
    for c in cams:
        d = dict(c)
        for k in ("viewmatrix", "projmatrix", "campos"):
            d[k] = t(c[k])
        cd.append(d)
    return g, cd


def run_ours(args, rank, world, local):
    import torch
    import torch.distributed as tdist
    import wheat_3dgs_b200
    from wheat_3dgs_b200 import dist as wdist
    from wheat_3dgs_b200 import ops, scene
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    sc, cams, (P, W, H, nv) = make_workload(args.workload)
    M = (SH_DEGREE + 1) ** 2
    g, cd = to_dev(sc, cams, dev)
    bg = torch.from_numpy(BG).to(dev)
    lg = scene.loss_gradients((3, H, W), seed=2)
    dc, dd, da = [torch.from_numpy(lg[k]).to(dev) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
    e = torch.Tensor([])
    # gradient bucket: symmetric memory + libwheatsplat's own all-reduce kernel when N > 1 (NCCL if unavailable)
    bucket, reduce_bucket, reduce_how = wdist.make_bucket(wdist.full_bucket_numel(P, M), dev)
    n_train = wdist.train_bucket_numel(P, M)

    # N > 1: views of similar cost (instance count R, from an untimed pass: preprocess + scan only) share a step,
    # because the step lasts as long as its slowest view (dist.balanced_schedule).  N = 1: plain round robin.
    schedule = None
    if world > 1:
        costs = [ops.count_instances(bg, g["means3D"], g["opacities"], g["scales"], g["rotations"], c["viewmatrix"],
                                     c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, g["shs"], SH_DEGREE, c["campos"])
                 for c in cd]
        schedule = wdist.balanced_schedule(costs, world, 2 * (args.steps + args.warmup) + 64)

    def view_of(i):
        if schedule is not None:
            return schedule[i % len(schedule)][rank]
        return (rank + i * world) % nv

    def fwd(c):
        return ops.rasterize_gaussians(bg, g["means3D"], e, g["opacities"], g["scales"], g["rotations"], 1.0, e,
                                       c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, g["shs"],
                                       SH_DEGREE, c["campos"], False, False)

    def bwd(c, fw):
        R, color, depth, alpha, radii, geom, binning, image = fw
        return ops.rasterize_gaussians_backward(bg, g["means3D"], radii, e, g["scales"], g["rotations"], 1.0, e,
                                                c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], dc, dd,
                                                da, g["shs"], SH_DEGREE, c["campos"], geom, R, binning, image, alpha,
                                                False, _bucket=bucket)

    def step(i, reduce=True):
        c = cd[view_of(i)]
        fw = fwd(c)
        bwd(c, fw)
        if reduce and world > 1:
            reduce_bucket(n_train)
        return fw

    def barrier():
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize()

    # ---- leg 1: device-resident throughput through the C-ABI operators
    for i in range(args.warmup):
        step(i)
    barrier()
    clk = ClockSampler(local)
    if rank == 0:   # one sampler per job: eight nvidia-smi pollers contend for the driver and slow every rank's launches
        clk.start()
        clk.wait_started()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = ops.launch_count()
    t0 = time.time()
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record()
    barrier()
    t1 = time.time()
    launches = ops.launch_count() - n0
    ms_total = wdist.max_over_ranks(e0.elapsed_time(e1), dev)
    clocks = clk.stop(t0, t1)

    # ---- leg 2: end to end through the drop-in module, host buffers in, loss out
    wheat_3dgs_b200.install()
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from wheat_3dgs_b200 import rasterizer as rz
    params = {k: v.clone().requires_grad_(True) for k, v in g.items()}
    means2D = torch.zeros_like(params["means3D"], requires_grad=True)
    n_img = min(nv, 8)
    gts = gt_images(n_img, H, W)
    cam_host = []
    for c in cams:
        h = torch.from_numpy(np.concatenate([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1), c["campos"]]))
        cam_host.append(h.pin_memory())
    h2d = 3 * H * W * 4 + 35 * 4
    rz.set_grad_bucket(bucket)

    pre = InputPrefetcher(dev, H, W, cam_host, gts, view_of)

    def e2e_step(i, last):
        v = view_of(i)
        sl = pre.take(i, last)
        cam, gt = sl["cam"], sl["gt"]
        rs = GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cams[v]["tanfovx"], tanfovy=cams[v]["tanfovy"], bg=bg,
            scale_modifier=1.0, viewmatrix=cam[0:16].view(4, 4), projmatrix=cam[16:32].view(4, 4),
            sh_degree=SH_DEGREE, campos=cam[32:35], prefiltered=False, debug=False)
        color, radii, depth, alpha = GaussianRasterizer(rs)(
            means3D=params["means3D"], means2D=means2D, opacities=params["opacities"], shs=params["shs"],
            scales=params["scales"], rotations=params["rotations"])
        loss = ((color - gt) ** 2).mean()
        loss.backward()
        pre.release(sl)
        if world > 1:
            reduce_bucket(n_train)
        res.push(loss)  # device -> host read of the step's result (ResultReader)
        for p_ in list(params.values()) + [means2D]:
            p_.grad = None

    res = ResultReader(torch)
    pre.prefetch(0)
    for i in range(args.warmup):
        e2e_step(i, last=False)
    res.flush()
    barrier()  # (the first timed step's inputs were prefetched during the last warm-up step, as every step's are)
    e0.record()
    for i in range(args.steps):
        e2e_step(args.warmup + i, last=(i == args.steps - 1))
    res.flush()
    e1.record()
    barrier()
    ms_e2e = wdist.max_over_ranks(e0.elapsed_time(e1), dev)
    rz.set_grad_bucket(None)

    # ---- leg 3: per-stage CUDA events (library-recorded, launching stream) + work counters
    stats = {"V": 0, "R": 0, "n_eval": 0, "n_contrib": 0}
    ops.prof_enable(True)
    torch.cuda.synchronize()
    ops.prof_read()
    nprof = min(args.steps, nv)
    for i in range(nprof):
        fw = step(args.warmup + i, reduce=False)
        stats["R"] += fw[0]
        stats["V"] += int((fw[4] > 0).sum())
        ne, nc = ops.count_pairs(W, H, fw[0], fw[6], fw[7])
        stats["n_eval"] += ne
        stats["n_contrib"] += nc
    torch.cuda.synchronize()
    prof = ops.prof_read()
    ops.prof_enable(False)
    for k in stats:
        stats[k] /= float(nprof)
    return dict(P=P, W=W, H=H, nv=nv, M=M, ms_total=ms_total, ms_e2e=ms_e2e, launches=launches, clocks=clocks,
                reduce_how=reduce_how,
                prof=prof, nprof=nprof, stats=stats, h2d=h2d, d2h=4)


# ----------------------------------------------------------------------------- reference arms
def run_reference_cuda(args, local):
    """The unmodified reference rasterizer (oracle/_ref/libref_dgr.so) on the same views, timed the
    same way.  Output / gradient zero-fills are ON (the reference op does them every call,
    DGR rasterize_points.cu:64-67,153-162); buffers are persistent (kinder than torch::empty)."""
    import torch
    from wheat_3dgs_b200 import scene
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref as refmod
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    sc, cams, (P, W, H, nv) = make_workload(args.workload)
    g, cd = to_dev(sc, cams, dev)
    bg = torch.from_numpy(BG).to(dev)
    lg = scene.loss_gradients((3, H, W), seed=2)
    dc, dd, da = [torch.from_numpy(lg[k]).to(dev) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
    rr = refmod.RefRasterizer("dgr")
    outs, grads = {}, {}

    def step(i, dcol=None):
        c = cd[i % nv]
        rr.forward(g, c, bg, SH_DEGREE, outputs=outs)
        rr.backward(dc if dcol is None else dcol, dd, da, grads=grads)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    clk = ClockSampler(local)
    clk.start()
    clk.wait_started()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record()
    torch.cuda.synchronize()
    t1 = time.time()
    ms_total = e0.elapsed_time(e1)
    clocks = clk.stop(t0, t1)

    # end to end: ground-truth image + camera from pinned host memory, loss back to the host
    n_img = min(nv, 8)
    gts = gt_images(n_img, H, W)
    cam_host = [torch.from_numpy(np.concatenate([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1),
                                                 c["campos"]])).pin_memory() for c in cams]

    pre = InputPrefetcher(dev, H, W, cam_host, gts, lambda i: i % nv)

    def e2e_step(i, last):
        v = i % nv
        sl = pre.take(i, last)
        cam, gt = sl["cam"], sl["gt"]
        c = dict(cams[v])
        c["viewmatrix"], c["projmatrix"], c["campos"] = cam[0:16], cam[16:32], cam[32:35]
        rr.forward(g, c, bg, SH_DEGREE, outputs=outs)
        diff = outs["color"] - gt
        loss = (diff ** 2).mean()
        dcol = diff * (2.0 / diff.numel())
        rr.backward(dcol, torch.zeros_like(dd), torch.zeros_like(da), grads=grads)
        pre.release(sl)
        res.push(loss)

    res = ResultReader(torch)
    pre.prefetch(0)
    for i in range(args.warmup):
        e2e_step(i, last=False)
    res.flush()
    torch.cuda.synchronize()
    e0.record()
    for i in range(args.steps):
        e2e_step(args.warmup + i, last=(i == args.steps - 1))
    res.flush()
    e1.record()
    torch.cuda.synchronize()
    ms_e2e = e0.elapsed_time(e1)
    return dict(P=P, W=W, H=H, nv=nv, ms_total=ms_total, ms_e2e=ms_e2e, clocks=clocks, h2d=3 * H * W * 4 + 35 * 4, d2h=4)


def run_cpu_oracle(workload, budget_s=12.0, max_views=3):
    """CPU oracle (a port: oracle/splat_oracle.c, OpenMP over tiles / Gaussians) fwd+bwd on a bounded
    sample of the workload's views.  Returns (views/s, threads, sample text)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from wheat_3dgs_b200 import scene
    sc, cams, (P, W, H, nv) = make_workload(workload)
    lg = scene.loss_gradients((3, H, W), seed=2)
    n, t_used = 0, 0.0
    while n < max_views and (n == 0 or t_used < budget_s):
        cam = cams[(17 + n) % nv]
        t0 = time.time()
        of = oracle.forward(sc, cam, BG, sh_degree=SH_DEGREE)
        oracle.backward(sc, cam, BG, of, lg["dL_dcolor"], lg["dL_ddepth"], lg["dL_dalpha"], sh_degree=SH_DEGREE)
        t_used += time.time() - t0
        n += 1
    return n / t_used, oracle.num_threads(), "%d view(s) of %s (P=%d, %dx%d) fwd+bwd, %.1f s" % (n, workload, P, W, H, t_used)


# ----------------------------------------------------------------------------- main
METRIC = "train views/s (rasterizer fwd+bwd per view, 1M Gaussians SH3, 2048x1536; 1000/value = fwd+bwd ms/view)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=144)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cpu"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    P, W, H, nv = WORKLOADS[args.workload]
    metric = METRIC if args.workload == "cfg2" else "train views/s (rasterizer fwd+bwd per view, %s)" % args.workload
    config = {"workload": "%s: synthetic plot_461-shaped scene, P=%d Gaussians SH3, %d overhead views %dx%d, fwd+bwd "
                          "per view" % (args.workload, P, nv, W, H),
              "views_per_step_per_gpu": 1, "parallelism": "views sharded x%d (views of similar instance count share a step), Gaussians "
              "replicated, 1 in-place SUM of 62*P fp32 per step" % world if world > 1 else "single GPU",
              "l2": "per-step inputs (236 MB Gaussians + 600 MB instance lists) exceed the 126 MB L2; no flush"}
    base = {"metric": metric, "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config}

    if args.impl != "ours":
        if rank != 0:
            return 0
        # the reference is single-GPU (utils/general_utils.py:133): under torchrun rank 0 alone runs it; n_gpus is the
        # launch's N so that the line pairs with this arm's line of the same N
        line = dict(base, impl="reference", n_gpus=world,
                    config=dict(config, parallelism="reference is single-GPU: one view per step on rank 0"
                                + (" (launched with %d ranks, the others idle)" % world if world > 1 else "")))
        if args.impl == "reference":
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import ref as refmod
            if not refmod.available("dgr"):
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_dgr.so not built "
                                  "(python oracle/build_ref.py needs /root/reference)"}))
                return 0
            r = run_reference_cuda(args, local)
            v = args.steps / (r["ms_total"] * 1e-3)
            line.update(value=v, ms_per_step=r["ms_total"] / args.steps, clocks=r["clocks"], gpu_launches=None,
                        reference_kind="unmodified reference CUDA rasterizer (diff-gaussian-rasterization @8829d14f) "
                        "built for sm_100, same B200; the reference has no CPU path",
                        e2e={"value": args.steps / (r["ms_e2e"] * 1e-3), "unit": "views/s",
                             "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"]})
            if not args.no_cpu_baseline:
                cv, cores, sample = run_cpu_oracle(args.workload)
                line["cpu_baseline"] = {"value": cv, "unit": "views/s", "cores": cores, "kind": "port", "sample": sample}
        else:
            cv, cores, sample = run_cpu_oracle(args.workload, budget_s=30.0, max_views=max(1, min(args.steps, 6)))
            line.update(value=cv, ms_per_step=1000.0 / cv, gpu_launches=0,
                        cpu_baseline={"value": cv, "unit": "views/s", "cores": cores, "kind": "port", "sample": sample},
                        e2e={"value": cv, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        print(json.dumps(line))
        return 0

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libwheatsplat has no CPU path")
    from wheat_3dgs_b200 import dist as wdist
    rank, world, local = wdist.init_from_env("nccl")
    r = run_ours(args, rank, world, local)
    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return 0
    peaks = load_peaks()
    st = r["stats"]
    T = ((W + 15) // 16) * ((H + 15) // 16)
    bits = max(1, int(np.ceil(np.log2(T + 1))))  # tile-id bits sorted per instance (depth-first binning)
    by, fl = stage_model(P, st["V"], st["R"], T, W * H, r["M"], (bits + 7) // 8, st["n_eval"], st["n_contrib"])
    stages = {}
    for name, (ms, calls) in r["prof"].items():
        per = ms / max(r["nprof"], 1)
        d = {"ms_per_view": per, "gbytes": by[name] / 1e9, "gbs": by[name] / 1e9 / (per * 1e-3) if per > 0 else None}
        if d["gbs"] is not None:
            d["hbm_frac"] = d["gbs"] / peaks["hbm_gbs"]
        if name in fl:
            d["tflops"] = fl[name] / 1e12 / (per * 1e-3) if per > 0 else None
        stages[name] = d
    top = max(stages, key=lambda k: stages[k]["ms_per_view"])
    traffic = None  # DRAM bytes per launch of the dominant kernel, from the committed ncu --set full capture
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            traffic = json.load(f)["bytes_per_launch"].get(top) if args.workload == "cfg2" else None
    except (OSError, ValueError, KeyError):
        pass
    sm_mhz = r["clocks"].get("sm_mhz") or peaks["sm_max_mhz"]
    if top in fl:
        peak = 148 * 128 * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12
        roof = {"kernel": top, "bound": "fp32", "achieved": stages[top]["tflops"], "peak": peak, "unit": "TFLOP/s",
                "frac": stages[top]["tflops"] / peak, "traffic": traffic,
                "peak_source": "148 SM x 128 lanes x 2 x %.0f MHz (clocks.max.sm); no measured fp32 figure in "
                               "MEASURED_PEAKS.json" % peaks["sm_max_mhz"],
                "frac_at_observed_clock": stages[top]["tflops"] / (148 * 128 * 2 * sm_mhz * 1e6 / 1e12),
                "hbm_gbs": stages[top]["gbs"], "hbm_frac": stages[top]["hbm_frac"]}
    else:
        roof = {"kernel": top, "bound": "hbm", "achieved": stages[top]["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": stages[top]["hbm_frac"], "traffic": traffic, "peak_source": peaks["source"]}
    value = world * args.steps / (r["ms_total"] * 1e-3)
    line = dict(base, impl="ours", value=value, ms_per_step=r["ms_total"] / args.steps, clocks=r["clocks"],
                gpu_launches=r["launches"], roofline=roof, stages=stages,
                work_per_view={k: st[k] for k in ("V", "R", "n_eval", "n_contrib")},
                e2e={"value": world * args.steps / (r["ms_e2e"] * 1e-3), "unit": "views/s",
                     "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"]})
    if world > 1:
        line["config"] = dict(config, all_reduce=r["reduce_how"])
    if world == 1 and not args.no_cpu_baseline:
        cv, cores, sample = run_cpu_oracle(args.workload)
        line["cpu_baseline"] = {"value": cv, "unit": "views/s", "cores": cores, "kind": "port", "sample": sample}
    print(json.dumps(line))
    if world > 1:
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
