#!/usr/bin/env python
"""bench.py -- the splat hot path on BASELINE.json's configs.

Default workload (`config.workload` = "cfg2"): synthetic plot_461-shaped scene, 1 M Gaussians, SH degree 3,
36 overhead views at 2048x1536 (wheat_3dgs_b200/scene.py, seeds 0/1/2).  One STEP = the rasterizer part of one
training iteration on one view: forward (preprocess, key emission, radix sort, tile ranges, blend with
depth+alpha) + backward (blend backward, preprocess backward) -- exactly what train_vanilla_3dgs.py:73,80
spends in the three submodules.  With N GPUs every rank takes its own views (weak scaling: N views per step,
Gaussians replicated) and the step ends with ONE in-place SUM of the 62*P-float gradient bucket.

  value      views/s over all ranks, inputs resident in HBM, C-ABI operator calls
  e2e        views/s through the drop-in GaussianRasterizer + autograd, with this step's ground-truth image and
             camera copied from pinned host memory and the loss read back inside the timed region
  roofline   the dominant stage, timed with CUDA events recorded by the library on the launching stream
  extra      BASELINE.json's other two metric parts measured in the same run at the same N:
             train_iters_s (full training iteration, workload "train") and render_mpix_s (orbit, "cfg4")
  exchange_check (N > 1)  the library's own all-reduce kernel against torch.distributed.all_reduce
  parity     this build's exact-tier hashes / image samples against tests/golden/fullsize.json (outputs of the
             unmodified reference CUDA rasterizer on a B200, tests/golden/make_golden_fullsize.py)
  cpu_baseline  the CPU oracle (oracle/splat_oracle.c, a port) on a bounded sample, rank 0, N=1

Other workloads (`--workload`): cfg1, cfg5 (6 M Gaussians, 3840x2160) = the same step; train = the whole
iteration of train_vanilla_3dgs.py:65-112; cfg3 = flashsplat label votes (multi-label vote pass over the views,
votes summed over ranks, multi_instance_opt; plus the run_3d_seg.py per-mask loop on a sample of seeds);
cfg4 = render_360 orbit frames, forward only, MPix/s.

`--impl reference` runs the UNMODIFIED reference CUDA rasterizers (oracle/_ref/libref_{dgr,fsr}.so, built from
/root/reference by oracle/build_ref.py) on the same GPU, same inputs, same timing: the reference has no CPU
implementation of this path, so the reference arm is its stock CUDA path.  `--impl reference-cpu` times the CPU
oracle port instead.  Neither touches libwheatsplat.so.
"""
import argparse
import hashlib
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: kind, P, W, H, views / frames
    "cfg2": dict(kind="fwdbwd", P=1000000, W=2048, H=1536, views=36),
    "cfg1": dict(kind="fwdbwd", P=50000, W=512, H=512, views=1),
    "cfg5": dict(kind="fwdbwd", P=6000000, W=3840, H=2160, views=144),
    "tiny": dict(kind="fwdbwd", P=20000, W=320, H=240, views=4),      # plumbing check only (tests)
    "train": dict(kind="train", P=1000000, W=2048, H=1536, views=36),
    "tinytrain": dict(kind="train", P=20000, W=320, H=240, views=4),
    "cfg4": dict(kind="orbit", P=1000000, W=1920, H=1080, views=200),
    "tiny4": dict(kind="orbit", P=20000, W=320, H=240, views=8),
    "cfg3": dict(kind="votes", P=1000000, W=2048, H=1536, views=36, labels=500),
    "tiny3": dict(kind="votes", P=20000, W=320, H=240, views=4, labels=12),
}
SH_DEGREE = 3
BG = np.array([0.0, 0.0, 0.0], np.float32)  # train_vanilla_3dgs.py:50 (black unless --white_background)
GOLDEN_FULLSIZE = os.path.join(ROOT, "tests", "golden", "fullsize.json")


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


def issue_active(kernel):
    """Issue-slot utilisation of `kernel` from the committed ncu --set full capture (profiles/r02_ncu_full_raw.csv):
    an issue-bound kernel's honest ceiling is issue slots, not FFMA lanes.  A profile figure, not a measurement of
    this run."""
    import csv
    try:
        with open(os.path.join(ROOT, "profiles", "r02_ncu_full_raw.csv")) as f:
            rows = list(csv.reader(f))
        hdr = rows[0]
        ki, ii = hdr.index("Kernel Name"), hdr.index("smsp__issue_active.avg.pct_of_peak_sustained_active")
        for r in rows[2:]:
            if kernel in r[ki]:
                return {"pct": float(r[ii]), "source": "profiles/r02_ncu_full_raw.csv (ncu --set full, config 2 view 17)"}
    except (OSError, ValueError, IndexError):
        pass
    return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": float(d["hbm_gbs"]), "sm_max_mhz": float(d.get("sm_max_mhz", 1965.0)), "source": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}


# ----------------------------------------------------------------------------- workload
def make_workload(name):
    from wheat_3dgs_b200 import orbit, scene
    wl = WORKLOADS[name]
    P, W, H, nv = wl["P"], wl["W"], wl["H"], wl["views"]
    sc = scene.synthetic_scene(P, seed=0, sh_degree=SH_DEGREE)
    if wl["kind"] == "orbit":
        # render_360.py:183 hands render_360 the first training view and scene.cameras_extent; frames are half
        # that view's resolution (utils/wheatgs_helper.py:277)
        og = scene.overhead_cameras(2 * W, 2 * H, seed=1)
        centers = np.stack([c["campos"] for c in og]).astype(np.float64)
        radius = 1.1 * float(np.linalg.norm(centers - centers.mean(axis=0), axis=1).max())  # getNerfppNorm
        centroid = sc["means3D"].astype(np.float64).mean(axis=0).tolist()                  # wheatgs_helper.py:276
        cams = orbit.orbit_views(og[0], radius, nv, elevation=45.0, centroid=centroid)
    elif nv == 144:
        cams = scene.overhead_cameras(W, H, n_x=24, n_y=6, seed=1)
    elif nv >= 36:
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


def cams_to_host(cams):
    import torch
    return [torch.from_numpy(np.concatenate([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1),
                                             c["campos"]])).pin_memory() for c in cams]


class InputPrefetcher:
    """The harness's input pipeline, shared by both arms (a data loader's prefetch).  Step i's camera (35
    floats) + per-view image (ground truth 3xHxW, or a 1xHxW label mask) travel from pinned host memory on a copy
    stream while step i-1 computes (two device slots); the compute stream waits for the copy's event.  Every
    step's H2D copy is issued and completes inside the timed region."""

    def __init__(self, dev, shape, cam_host, imgs, view_of):
        import torch
        self.torch, self.cam_host, self.imgs, self.view_of = torch, cam_host, imgs, view_of
        self.stream = torch.cuda.Stream(device=dev)
        self.slots = [{"cam": torch.empty(35, dtype=torch.float32, device=dev),
                       "gt": torch.empty(shape, dtype=torch.float32, device=dev) if shape else None,
                       "ready": torch.cuda.Event(), "free": torch.cuda.Event()} for _ in range(2)]
        for sl in self.slots:
            sl["free"].record()
        self.copies = []   # (start event, end event, bytes) of every step's upload, on the copy stream

    def prefetch(self, i):
        v = self.view_of(i)
        sl = self.slots[i & 1]
        with self.torch.cuda.stream(self.stream):
            self.stream.wait_event(sl["free"])  # the step that last read this slot has finished with it
            t0 = self.torch.cuda.Event(enable_timing=True)
            t0.record(self.stream)
            sl["cam"].copy_(self.cam_host[v], non_blocking=True)
            nbytes = sl["cam"].numel() * 4
            if sl["gt"] is not None:
                sl["gt"].copy_(self.imgs[v % len(self.imgs)], non_blocking=True)
                nbytes += sl["gt"].numel() * 4
            sl["ready"].record(self.stream)
            t1 = self.torch.cuda.Event(enable_timing=True)
            t1.record(self.stream)
            self.copies.append((t0, t1, nbytes))

    def copy_stats(self, last_n):
        """Duration of the last `last_n` uploads as the copy stream saw them (CUDA events either side of the copies,
        while the previous step's kernels run beside them): mean / max milliseconds and the mean rate in GB/s.  Call
        after the device has drained."""
        c = self.copies[-last_n:] if last_n > 0 else []
        if not c:
            return None
        ms = [a.elapsed_time(b) for a, b, _ in c]
        nb = sum(n for _, _, n in c) / len(c)
        mean = sum(ms) / len(ms)
        return {"h2d_ms_mean": mean, "h2d_ms_max": max(ms), "h2d_gbs": nb / (mean * 1e-3) / 1e9 if mean > 0 else None,
                "copies": len(c)}

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
    """The harness's result pipeline, shared by both arms.  Every step's result scalar (4 bytes) is copied to pinned
    host memory behind the step's kernels and READ on the host one step later, once the next step has been queued
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


def stage_model(P, V, R, T, HW, M, npass, n_eval, n_contrib, Rf=None):
    """Algorithmic HBM bytes / flops per view of every stage (DESIGN.md section 3 derives them)."""
    c = 12 * M  # SH bytes per Gaussian
    Rf = R if Rf is None else Rf
    by = {
        # depth-first binning (default): preprocess also writes the (depth bits, id) pair; "sort" =
        # the depth sort of P 8-byte pairs + the tile sort of R 8-byte pairs (npass passes)
        "preprocess": P * (44 + 8 + 8) + V * (c + 65),
        "scan": 8 * ((P + 255) // 256),
        "emit_keys": 12 * P + 8 * ((P + 255) // 256) + 40 * V + 8 * R,
        # (the tile sort's digit histograms are counted by the emit pass: no read of the R keys for them)
        "sort": (4 * P + 4 * 16 * P) + npass * 16 * R,
        # identify_ranges reads the sorted tile keys only; the blend kernels read the value list and gather the
        # 64-byte rows of the instances whose quadrant mask is not empty (Rf of them) from the per-Gaussian table
        "ranges_pack": 4 * R + 8 * T,
        "render_fwd": 4 * R + 64 * Rf + 24 * HW,
        "render_bwd": 4 * R + 64 * Rf + 32 * HW + 40 * V,
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


class Job:
    """Rank / device bookkeeping and the timing bracket of the contract: barrier + synchronize on both sides, CUDA
    events on the launching stream, max over ranks."""

    def __init__(self, rank, world, local):
        import torch
        self.torch, self.rank, self.world, self.local = torch, rank, world, local
        self.dev = torch.device("cuda", local)
        torch.cuda.set_device(self.dev)

    def barrier(self):
        if self.world > 1:
            self.torch.distributed.barrier()
        self.torch.cuda.synchronize()

    def timed(self, step, steps, warmup, before=None, after=None):
        """ms over `steps` calls of step(i) after `warmup` untimed ones (max over ranks)."""
        from wheat_3dgs_b200 import dist as wdist
        torch = self.torch
        for i in range(warmup):
            step(i)
        if after is not None:
            after()
        self.barrier()
        if before is not None:
            before()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            step(warmup + i)
        if after is not None:
            after()
        e1.record()
        self.barrier()
        return wdist.max_over_ranks(e0.elapsed_time(e1), self.dev)


def sha(t):
    """sha256 of a tensor's bytes (host side)."""
    return hashlib.sha256(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes()).hexdigest()


def load_golden(key):
    try:
        with open(GOLDEN_FULLSIZE) as f:
            return json.load(f).get(key)
    except (OSError, ValueError):
        return None


GOLDEN_SAMPLES = 8192  # image / gradient samples kept per tensor in the full-size goldens (evenly spaced)


def sample(t):
    """GOLDEN_SAMPLES evenly spaced elements of a tensor, as float32 numpy."""
    flat = t.detach().reshape(-1)
    n = min(flat.numel(), GOLDEN_SAMPLES)
    if n == 0:
        return np.zeros((0,), np.float32)
    step = flat.numel() // n
    return flat[:: step][:n].float().cpu().numpy()


def load_golden_samples(prefix):
    try:
        z = np.load(os.path.join(ROOT, "tests", "golden", "fullsize.npz"))
        return {k[len(prefix) + 1:]: z[k] for k in z.files if k.startswith(prefix + "/")}
    except (OSError, ValueError):
        return {}


def fwd_signature(fw, P, W, H):
    """Exact-tier hashes and image samples of one finished forward (what tests/golden/fullsize.json stores)."""
    from wheat_3dgs_b200 import ops
    R, color, depth, alpha, radii, geom, binning, image = fw
    sig = {"num_rendered": int(R), "radii": sha(radii),
           "point_list": sha(ops.debug_fetch("point_list", P, W, H, R, geom, binning, image)),
           "ranges": sha(ops.debug_fetch("ranges", P, W, H, R, geom, binning, image)),
           "n_contrib": sha(ops.debug_fetch("n_contrib", P, W, H, R, geom, binning, image))}
    samples = {k: sample(t) for k, t in (("color", color), ("depth", depth), ("alpha", alpha))}
    return sig, samples


def compare_signature(gold, sig, samples, grads=None, name=None):
    """-> parity dict for the bench line."""
    if gold is None:
        return {"checked": False, "why": "no entry in tests/golden/fullsize.json (run tests/golden/make_golden_fullsize.py on a GPU box)"}
    out = {"checked": True, "against": "unmodified reference CUDA rasterizer outputs (tests/golden/fullsize.json)",
           "view": gold.get("view")}
    exact = True
    for k in ("num_rendered", "radii", "point_list", "ranges", "n_contrib"):
        ok = gold[k] == sig[k]
        out[k + "_equal"] = bool(ok)
        exact = exact and ok
    worst = 0.0
    gs = load_golden_samples(name)
    for k, v in samples.items():
        worst = max(worst, float(np.abs(v - gs[k]).max()) if v.size else 0.0)
    out["image_max_abs_err"] = worst
    ok = exact and worst <= 1e-4
    if grads is not None and all(k in gs for k in grads):
        rel = {}
        for k, v in grads.items():
            ref = np.asarray(gs[k], np.float64)
            n = np.linalg.norm(ref)
            rel[k] = float(np.linalg.norm(v.astype(np.float64) - ref) / n) if n > 0 else 0.0
        out["grad_norm_rel_err"] = rel
        ok = ok and all(r < 1e-3 for r in rel.values())
    out["ok"] = bool(ok)
    return out


GRAD_KEYS = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations")


def exchange_check(job, bucket, reduce_bucket, n):
    """The library's own all-reduce kernel against torch.distributed.all_reduce on a rank-seeded pattern."""
    import torch
    import torch.distributed as tdist
    g = torch.Generator(device=job.dev).manual_seed(4242 + job.rank)
    keep = bucket.clone()
    bucket.copy_(torch.randn(bucket.numel(), generator=g, device=job.dev, dtype=torch.float32))
    # a few huge / tiny magnitudes and exact zeros so that order-dependent rounding would show
    bucket[:: 1009] *= 1e6
    bucket[5:: 2003] = 0.0
    ref = bucket.clone()
    job.barrier()
    reduce_bucket(n)
    tdist.all_reduce(ref[:n], op=tdist.ReduceOp.SUM)
    torch.cuda.synchronize()
    scale = float(ref[:n].abs().max())
    err = float((bucket[:n] - ref[:n]).abs().max()) / scale
    tail_untouched = bool(torch.equal(bucket[(n + 3) // 4 * 4:], ref[(n + 3) // 4 * 4:]))
    words = bucket[:n].view(torch.int32).to(torch.int64)
    chk = torch.stack([words.sum(), (words * (torch.arange(n, device=job.dev) % 8191 + 1)).sum()])
    allc = [torch.zeros_like(chk) for _ in range(job.world)]
    tdist.all_gather(allc, chk)
    same = all(bool(torch.equal(c, allc[0])) for c in allc)
    bucket.copy_(keep)
    job.barrier()
    return {"numel": int(n), "max_rel_err_vs_nccl": err, "identical_across_ranks": same,
            "tail_untouched": tail_untouched, "ok": bool(err < 1e-5 and same and tail_untouched)}


# ----------------------------------------------------------------------------- fwd+bwd (cfg1 / cfg2 / cfg5)
def run_fwdbwd(args, job):
    import torch
    import wheat_3dgs_b200
    from wheat_3dgs_b200 import dist as wdist
    from wheat_3dgs_b200 import ops, scene
    rank, world, dev = job.rank, job.world, job.dev
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

    # ---- parity gate (untimed): the golden view against the reference's outputs, hashes and samples
    gold = load_golden(args.workload)
    parity = None
    if rank == 0:
        if gold is None:
            parity = compare_signature(None, None, None)
        else:
            c = cd[gold["view"]]
            fw = fwd(c)
            sig, samples = fwd_signature(fw, P, W, H)
            grads = bwd(c, fw)
            gsm = {k: sample(t) for k, t in zip(GRAD_KEYS, grads)}
            parity = compare_signature(gold, sig, samples, gsm, name=args.workload)
            del fw, grads
    xchk = exchange_check(job, bucket, reduce_bucket, n_train) if world > 1 else None

    # ---- leg 1: device-resident throughput through the C-ABI operators
    for i in range(args.warmup):
        step(i)
    job.barrier()
    clk = ClockSampler(job.local)
    if rank == 0:   # one sampler per job: eight nvidia-smi pollers contend for the driver and slow every rank's launches
        clk.start()
        clk.wait_started()
    job.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = ops.launch_count()
    t0 = time.time()
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record()
    job.barrier()
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
    gts = gt_images(min(nv, 8), H, W)
    cam_host = cams_to_host(cams)
    h2d = 3 * H * W * 4 + 35 * 4
    rz.set_grad_bucket(bucket)
    pre = InputPrefetcher(dev, (3, H, W), cam_host, gts, view_of)

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
    job.barrier()  # (the first timed step's inputs were prefetched during the last warm-up step, as every step's are)
    e0.record()
    for i in range(args.steps):
        e2e_step(args.warmup + i, last=(i == args.steps - 1))
    res.flush()
    e1.record()
    job.barrier()
    ms_e2e = wdist.max_over_ranks(e0.elapsed_time(e1), dev)
    torch.cuda.synchronize()
    h2d_stats = pre.copy_stats(args.steps)   # this rank's uploads of the timed steps
    rz.set_grad_bucket(None)
    del params, means2D, pre, gts

    # ---- leg 3: per-stage CUDA events (library-recorded, launching stream) + work counters
    stats = {"V": 0, "R": 0, "n_eval": 0, "n_contrib": 0, "quad_pairs": 0, "instances_never_fetched": 0}
    ops.prof_enable(True)
    torch.cuda.synchronize()
    ops.prof_read()
    nprof = min(args.steps, nv)
    for i in range(nprof):
        fw = step(args.warmup + i, reduce=False)
        stats["R"] += fw[0]
        stats["V"] += int((fw[4] > 0).sum())
        ne, nc, nq, nz = ops.count_pairs(P, W, H, fw[0], fw[5], fw[6], fw[7])
        stats["n_eval"] += ne
        stats["n_contrib"] += nc
        stats["quad_pairs"] += nq
        stats["instances_never_fetched"] += nz
    torch.cuda.synchronize()
    prof = ops.prof_read()
    ops.prof_enable(False)
    for k in stats:
        stats[k] /= float(nprof)

    # ---- leg 4 (N > 1): where the step's time beyond one view goes -- the exchange kernel alone (barriers
    # included, ranks already aligned) against the whole step
    xsplit = None
    if world > 1:
        job.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        reduce_bucket(n_train)
        job.barrier()
        ev[0].record()
        for _ in range(10):
            reduce_bucket(n_train)
        ev[1].record()
        job.barrier()
        xsplit = {"exchange_ms_aligned": wdist.max_over_ranks(ev[0].elapsed_time(ev[1]) / 10, dev)}
    del fw
    fma = {"ffma2_tflops": ops.fma_peak(True, dev), "ffma_tflops": ops.fma_peak(False, dev)} if rank == 0 else None
    return dict(P=P, W=W, H=H, nv=nv, M=M, ms_total=ms_total, ms_e2e=ms_e2e, launches=launches, clocks=clocks,
                reduce_how=reduce_how, prof=prof, nprof=nprof, stats=stats, h2d=h2d, d2h=4, parity=parity,
                exchange_check=xchk, exchange_split=xsplit, fma=fma, h2d_stats=h2d_stats)


# ----------------------------------------------------------------------------- full training iteration
def run_train(args, job, workload, steps, warmup, with_e2e=True):
    """train_vanilla_3dgs.py:65-112 on the fused stack (wheat_3dgs_b200/train.py): raw parameters -> forward_raw ->
    fused L1+SSIM -> backward -> statistics -> exchange + Adam.  value = views/s = N x iterations/s."""
    import torch
    from wheat_3dgs_b200 import dist as wdist
    from wheat_3dgs_b200 import ops, train
    rank, world, dev = job.rank, job.world, job.dev
    sc, cams, (P, W, H, nv) = make_workload(workload)
    _, cd = to_dev({k: sc[k][:1] for k in ("means3D", "scales", "rotations", "opacities", "shs")}, cams, dev)
    bg = torch.from_numpy(BG).to(dev)
    tr = train.Trainer(train.raw_params_from_scene(sc, dev), SH_DEGREE, dev, world=world)
    gts_host = gt_images(min(nv, 4), H, W)
    gts = [t.to(dev) for t in gts_host]
    view_of = lambda i: (rank + i * world) % nv
    losses = []

    # parity gate: the first iterations' losses against the reference stack's (reference rasterizer + torch loss +
    # torch.optim.Adam from the same initial parameters; tests/golden/fullsize.json), single GPU semantics
    gold = load_golden(workload)
    parity = {"checked": False, "why": "no entry in tests/golden/fullsize.json"} if gold is None else None
    if gold is not None and world == 1:
        chk = train.Trainer(train.raw_params_from_scene(sc, dev), SH_DEGREE, dev, world=1)
        mine = [float(chk.iteration(cd[i % nv], gts[i % len(gts)], bg)) for i in range(len(gold["losses"]))]
        rel = [abs(a - b) / abs(b) for a, b in zip(mine, gold["losses"])]
        parity = {"checked": True, "against": "reference stack losses of the first iterations (tests/golden/fullsize.json)",
                  "losses": mine, "reference_losses": gold["losses"], "max_rel_err": max(rel), "ok": bool(max(rel) < 1e-4)}
        del chk
    elif gold is not None:
        parity = {"checked": False, "why": "the loss golden is single-GPU semantics (one view per step); checked at N=1"}

    def step(i):
        losses.append(tr.iteration(cd[view_of(i)], gts[i % len(gts)], bg))

    n0 = ops.launch_count()
    ms = job.timed(step, steps, warmup)
    launches = ops.launch_count() - n0
    out = dict(P=P, W=W, H=H, nv=nv, ms_total=ms, how=tr.how, launches=launches, parity=parity,
               final_loss=float(losses[-1]))
    if with_e2e:
        cam_host = cams_to_host(cams)
        pre = InputPrefetcher(dev, (3, H, W), cam_host, gts_host, view_of)
        res = ResultReader(torch)

        def e2e_step(i):
            v = view_of(i)
            sl = pre.take(i, False)
            cam = dict(cams[v], viewmatrix=sl["cam"][0:16].view(4, 4), projmatrix=sl["cam"][16:32].view(4, 4),
                       campos=sl["cam"][32:35])
            loss = tr.iteration(cam, sl["gt"], bg)
            pre.release(sl)
            res.push(loss)

        pre.prefetch(0)
        out["ms_e2e"] = job.timed(e2e_step, steps, warmup, after=res.flush)
        out["h2d"], out["d2h"] = 3 * H * W * 4 + 35 * 4, 4
    del tr
    return out


# ----------------------------------------------------------------------------- orbit frames (cfg4)
def run_orbit(args, job, workload, steps, warmup, with_e2e=True):
    """utils/wheatgs_helper.py:282-300: frames rank::world of the orbit, forward only with depth + alpha.
    value = MPix/s over all ranks, frames left on the rendering GPU; e2e = every frame (color) copied to pinned
    host memory by its own rank (each rank writes its own PNGs; the reference does rendering.cpu(), :291);
    gathered = frames collected on rank 0's GPU with one NCCL gather per step (dist / SURVEY 8e)."""
    import torch
    import torch.distributed as tdist
    from wheat_3dgs_b200 import ops, orbit
    rank, world, dev = job.rank, job.world, job.dev
    sc, cams, (P, W, H, nv) = make_workload(workload)
    g, cd = to_dev(sc, cams, dev)
    bg = torch.ones(3, device=dev)   # render_360.py:146: white background
    frame_of = lambda i: (rank + i * world) % nv
    parity = None
    gold = load_golden(workload)
    if rank == 0:
        if gold is None:
            parity = compare_signature(None, None, None)
        else:
            c = cd[gold["view"]]
            fw = ops.rasterize_gaussians(bg, g["means3D"], orbit._E, g["opacities"], g["scales"], g["rotations"], 1.0,
                                         orbit._E, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W,
                                         g["shs"], SH_DEGREE, c["campos"], False, False)
            sig, samples = fwd_signature(fw, P, W, H)
            parity = compare_signature(gold, sig, samples, name=workload)
            del fw
    # frames are independent: they go on three streams in turn (orbit.StreamRing, as orbit.render_360 does), so the
    # short binning kernels of one frame run under the blend kernel of another
    ring = orbit.StreamRing(int(os.environ.get("WS_ORBIT_STREAMS", "3")), dev)
    keep = [None] * ring.n

    def step(i):
        with ring.stream(i):
            keep[i % ring.n] = orbit.render_frame(g, cd[frame_of(i)], bg, SH_DEGREE)

    n0 = ops.launch_count()
    ms = job.timed(step, steps, warmup, before=ring.fork, after=ring.join)
    launches = ops.launch_count() - n0
    out = dict(P=P, W=W, H=H, nv=nv, ms_total=ms, launches=launches, parity=parity)
    if with_e2e:
        cam_host = cams_to_host(cams)
        pre = InputPrefetcher(dev, None, cam_host, None, frame_of)
        host = [torch.empty((3, H, W), dtype=torch.float32).pin_memory() for _ in range(2)]
        evs = [torch.cuda.Event(), torch.cuda.Event()]
        copy_stream = torch.cuda.Stream(device=dev)
        done = torch.cuda.Event()
        nsteps = [0]

        def e2e_step(i):
            f = frame_of(i)
            with ring.stream(i):
                sl = pre.take(i, False)
                cam = dict(cams[f], viewmatrix=sl["cam"][0:16].view(4, 4), projmatrix=sl["cam"][16:32].view(4, 4),
                           campos=sl["cam"][32:35])
                color = orbit.render_frame(g, cam, bg, SH_DEGREE)[0]
                pre.release(sl)
                done.record()
            k = nsteps[0] & 1
            nsteps[0] += 1
            evs[k].synchronize()            # the host buffer's previous frame has landed (it was "encoded")
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(done)
                host[k].copy_(color, non_blocking=True)
                color.record_stream(copy_stream)
                evs[k].record(copy_stream)

        def drain():
            for e_ in evs:
                e_.synchronize()
            torch.cuda.current_stream().wait_stream(copy_stream)
            ring.join()

        for e_ in evs:
            e_.record()
        pre.prefetch(0)
        out["ms_e2e"] = job.timed(e2e_step, steps, warmup, after=drain)
        out["h2d"], out["d2h"] = 35 * 4, 3 * H * W * 4
        if world > 1:
            gl = [torch.empty((3, H, W), dtype=torch.float32, device=dev) for _ in range(world)] if rank == 0 else None

            def gather_step(i):
                color = orbit.render_frame(g, cd[frame_of(i)], bg, SH_DEGREE)[0]
                tdist.gather(color, gl, dst=0)

            out["ms_gathered"] = job.timed(gather_step, steps, warmup)
    return out


# ----------------------------------------------------------------------------- label votes (cfg3)
def make_masks(sc, cams, nobj, views):
    from wheat_3dgs_b200 import scene
    return {v: scene.head_masks(sc, cams[v], n_labels=nobj, solid=True) for v in views}


def run_votes(args, job, workload, steps, warmup):
    """BASELINE config 3 in its two forms.
    (i)  multi-label vote pass (render_360.py:168-181 / eval_wheatgs.py:106-114): views sharded over the ranks, each a
         flashsplat forward with that view's ~500-id mask accumulating into ONE persistent used_count[501, P] per
         rank, summed over ranks by the exchange kernel, then multi_instance_opt.  value = views/s.
    (ii) run_3d_seg.py:243-315 per seed mask (segment.HeadSegmenter): obj_num = 1 votes, labels, find_match over the
         other views at IoU 0.6, growing loop; seeds sharded over the ranks (independent objects).  extra.seeds_s."""
    import torch
    from wheat_3dgs_b200 import dist as wdist
    from wheat_3dgs_b200 import ops, segment
    rank, world, dev = job.rank, job.world, job.dev
    wl = WORKLOADS[workload]
    nobj = wl["labels"]
    sc, cams, (P, W, H, nv) = make_workload(workload)
    g, cd = to_dev(sc, cams, dev)
    bg = torch.from_numpy(BG).to(dev)
    e = torch.Tensor([])
    my_views = list(range(rank, nv, world))
    masks_host = {v: torch.from_numpy(m).pin_memory() for v, m in make_masks(sc, cams, nobj, range(nv)).items()}
    masks = {v: m.to(dev) for v, m in masks_host.items()}
    rows = nobj + 1
    n_votes = rows * P
    votes, reduce_votes, how = wdist.make_bucket(n_votes, dev)
    counts = votes[:n_votes].view(rows, P)

    def vote_view(v, mask=None, cam=None):
        c = cd[v] if cam is None else cam
        return ops.flash_rasterize_gaussians(masks[v] if mask is None else mask, None, bg, g["means3D"], e,
                                             g["opacities"], g["scales"], g["rotations"], 1.0, e, c["viewmatrix"],
                                             c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, g["shs"], SH_DEGREE,
                                             c["campos"], False, nobj, False, used_count_out=counts)

    # ---- parity gate: conservation (the votes of a view sum to its alpha over labelled pixels) on this rank's first
    # view, and -- N = 1 -- the label row sums after the golden's views against the reference FSR's
    counts.zero_()
    out_v = vote_view(my_views[0])
    cons = abs(float(counts.double().sum()) - float(out_v[3].double().sum())) / max(float(out_v[3].double().sum()), 1e-9)
    parity = {"votes_vs_alpha_rel_err": cons, "ok": bool(cons < 1e-4)}
    gold = load_golden(workload)
    if gold is None:
        parity.update(checked=False, why="no entry in tests/golden/fullsize.json")
    elif world == 1:
        counts.zero_()
        for v in gold["views"]:
            vote_view(v)
        lab = ops.multi_instance_opt(counts)
        rowsum = lab.sum(dim=1).cpu().numpy().astype(np.int64)
        ref = np.asarray(gold["label_row_sums"], np.int64)
        diff = int(np.abs(rowsum - ref).sum())
        parity.update(checked=True, against="reference FSR votes + reference multi_instance_opt (tests/golden/fullsize.json)",
                      views=gold["views"], labelled=int(rowsum[1:].sum()), reference_labelled=int(ref[1:].sum()),
                      row_sum_l1_diff=diff, reference_near_ties=gold.get("near_ties"),
                      ok=bool(parity["ok"] and diff <= max(4, 2 * int(gold.get("near_ties", 0)))))
        del lab
    del out_v

    # ---- (i) vote pass: one step = one view per rank
    # (views are independent and the votes are float atomics into one accumulator: the views go on three streams in
    # turn, orbit.StreamRing, so one view's binning kernels run under another's blend)
    from wheat_3dgs_b200 import orbit
    ring = orbit.StreamRing(int(os.environ.get("WS_ORBIT_STREAMS", "3")), dev)
    view_of = lambda i: my_views[i % len(my_views)]

    def vote_step(i):
        with ring.stream(i):
            vote_view(view_of(i))

    def fresh():
        counts.zero_()
        ring.fork()

    n0 = ops.launch_count()
    ms = job.timed(vote_step, steps, warmup, before=fresh, after=ring.join)
    launches = ops.launch_count() - n0
    # the exchange + label assignment that end the pass (once per pass, timed separately)
    job.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record()
    if world > 1:
        reduce_votes(n_votes)
    ev[1].record()
    labels = ops.multi_instance_opt(counts)
    ev[2].record()
    job.barrier()
    ms_reduce = wdist.max_over_ranks(ev[0].elapsed_time(ev[1]), dev)
    ms_labels = wdist.max_over_ranks(ev[1].elapsed_time(ev[2]), dev)
    labelled = int(labels[1:].any(dim=0).sum())
    del labels

    # e2e: the view's id mask (HxW fp32) + camera from pinned host memory every step, num_rendered back to the host
    cam_host = cams_to_host(cams)
    pre = InputPrefetcher(dev, (H, W), cam_host, [masks_host[v] for v in range(nv)], view_of)

    def e2e_step(i):
        v = view_of(i)
        with ring.stream(i):
            sl = pre.take(i, False)
            cam = dict(cams[v], viewmatrix=sl["cam"][0:16].view(4, 4), projmatrix=sl["cam"][16:32].view(4, 4),
                       campos=sl["cam"][32:35])
            vote_view(v, mask=sl["gt"], cam=cam)   # num_rendered comes back to the host inside the call
            pre.release(sl)

    pre.prefetch(0)
    ms_e2e = job.timed(e2e_step, steps, warmup, before=fresh, after=ring.join)

    # ---- (ii) run_3d_seg per-seed loop on a sample of seeds (this rank's share)
    seg = segment.HeadSegmenter(g, cd, masks, bg, SH_DEGREE, iou_threshold=0.6)
    seed_view = my_views[0]
    ids = torch.unique(masks[seed_view]).cpu().numpy().astype(np.int64)
    ids = [int(k) for k in ids if k > 0]
    n_seeds = min(args.seeds, len(ids))
    seeds = ids[:: max(1, len(ids) // max(n_seeds, 1))][:n_seeds]
    results = []
    if seeds:
        shared = seg.view_counts(seed_view, nobj)
        torch.cuda.synchronize()
        t0 = time.time()
        r0 = seg.renders
        for k in seeds:
            results.append(seg.segment_seed(seed_view, k, shared=shared))
        torch.cuda.synchronize()
        dt = time.time() - t0
        renders = seg.renders - r0
        del shared
    else:
        dt, renders = 0.0, 0
    dt = wdist.max_over_ranks(dt, dev)
    found = [r for r in results if r is not None]
    seg_out = {"seeds": len(seeds) * world, "seconds": dt, "seeds_s": (len(seeds) * world / dt) if dt > 0 else None,
               "renders_rank0": renders, "matched_views_mean_rank0": float(np.mean([len(r["matched"]) for r in found])) if found else 0.0,
               "gaussians_mean_rank0": float(np.mean([int(r["used_mask"].sum()) for r in found])) if found else 0.0,
               "views_matched_against": nv - 1}
    return dict(P=P, W=W, H=H, nv=nv, ms_total=ms, ms_e2e=ms_e2e, launches=launches, parity=parity, how=how,
                ms_reduce=ms_reduce, ms_labels=ms_labels, labelled=labelled, seg=seg_out, h2d=H * W * 4 + 35 * 4, d2h=4,
                rows=rows)


# ----------------------------------------------------------------------------- reference arms
def _ref_inputs(workload, local):
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref as refmod
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    sc, cams, dims = make_workload(workload)
    g, cd = to_dev(sc, cams, dev)
    return torch, refmod, dev, sc, cams, g, cd, dims


def _timed_single(torch, step, steps, warmup, after=None):
    for i in range(warmup):
        step(i)
    if after is not None:
        after()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(warmup + i)
    if after is not None:
        after()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def run_reference_fwdbwd(args, local):
    """The unmodified reference rasterizer (oracle/_ref/libref_dgr.so) on the same views, timed the
    same way.  Output / gradient zero-fills are ON (the reference op does them every call,
    DGR rasterize_points.cu:64-67,153-162); buffers are persistent (kinder than torch::empty)."""
    from wheat_3dgs_b200 import scene
    torch, refmod, dev, sc, cams, g, cd, (P, W, H, nv) = _ref_inputs(args.workload, local)
    bg = torch.from_numpy(BG).to(dev)
    lg = scene.loss_gradients((3, H, W), seed=2)
    dc, dd, da = [torch.from_numpy(lg[k]).to(dev) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
    rr = refmod.RefRasterizer("dgr")
    outs, grads = {}, {}

    def step(i):
        rr.forward(g, cd[i % nv], bg, SH_DEGREE, outputs=outs)
        rr.backward(dc, dd, da, grads=grads)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    clk = ClockSampler(local)
    clk.start()
    clk.wait_started()
    t0 = time.time()
    ms_total = _timed_single(torch, step, args.steps, 0)
    t1 = time.time()
    clocks = clk.stop(t0, t1)

    # end to end: ground-truth image + camera from pinned host memory, loss back to the host
    gts = gt_images(min(nv, 8), H, W)
    pre = InputPrefetcher(dev, (3, H, W), cams_to_host(cams), gts, lambda i: i % nv)
    zd, za = torch.zeros_like(dd), torch.zeros_like(da)

    def e2e_step(i):
        v = i % nv
        sl = pre.take(i, False)
        cam, gt = sl["cam"], sl["gt"]
        c = dict(cams[v])
        c["viewmatrix"], c["projmatrix"], c["campos"] = cam[0:16], cam[16:32], cam[32:35]
        rr.forward(g, c, bg, SH_DEGREE, outputs=outs)
        diff = outs["color"] - gt
        loss = (diff ** 2).mean()
        dcol = diff * (2.0 / diff.numel())
        rr.backward(dcol, zd, za, grads=grads)
        pre.release(sl)
        res.push(loss)

    res = ResultReader(torch)
    pre.prefetch(0)
    ms_e2e = _timed_single(torch, e2e_step, args.steps, args.warmup, after=res.flush)
    return dict(P=P, W=W, H=H, nv=nv, ms_total=ms_total, ms_e2e=ms_e2e, clocks=clocks, h2d=3 * H * W * 4 + 35 * 4, d2h=4)


def reference_trainer(local, workload):
    """The reference's training iteration on the same GPU: torch getters -> UNMODIFIED reference rasterizer behind an
    autograd Function (with its zero fills) -> utils/loss_utils.py semantics in torch (5 grouped conv2d) -> backward
    -> the reference's masked-index statistics -> torch.optim.Adam."""
    import torch.nn.functional as F
    from wheat_3dgs_b200 import train
    torch, refmod, dev, sc, cams, g, cd, (P, W, H, nv) = _ref_inputs(workload, local)
    del g
    bg = torch.from_numpy(BG).to(dev)
    rr = refmod.RefRasterizer("dgr")

    class RefRaster(torch.autograd.Function):
        @staticmethod
        def forward(ctx, cam, means3D, means2D, shs, opacities, scales, rotations):
            gg = dict(means3D=means3D.detach(), shs=shs.detach(), opacities=opacities.detach(), scales=scales.detach(),
                      rotations=rotations.detach())
            o = rr.forward(gg, cam, bg, SH_DEGREE, outputs={})
            ctx.mark_non_differentiable(o["radii"])
            return o["color"], o["radii"], o["depth"], o["alpha"]

        @staticmethod
        def backward(ctx, gc, gr, gd, ga):
            o = rr.last["out"]
            z = lambda x, r: torch.zeros_like(r) if x is None else x.contiguous()
            gr_ = rr.backward(z(gc, o["color"]), z(gd, o["depth"]), z(ga, o["alpha"]), grads={})
            return (None, gr_["dL_dmeans3D"], gr_["dL_dmeans2D"], gr_["dL_dsh"], gr_["dL_dopacity"], gr_["dL_dscales"],
                    gr_["dL_drotations"])

    g1 = torch.tensor([math.exp(-(i - 5) ** 2 / 4.5) for i in range(11)])
    g1 = (g1 / g1.sum()).unsqueeze(1)
    win = g1.mm(g1.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().to(dev)

    def torch_loss(a, b, lam=0.2):   # utils/loss_utils.py:17-62 as train_vanilla_3dgs.py:77-79 combines them
        mu1, mu2 = F.conv2d(a, win, padding=5, groups=3), F.conv2d(b, win, padding=5, groups=3)
        s1 = F.conv2d(a * a, win, padding=5, groups=3) - mu1 * mu1
        s2 = F.conv2d(b * b, win, padding=5, groups=3) - mu2 * mu2
        s12 = F.conv2d(a * b, win, padding=5, groups=3) - mu1 * mu2
        m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
        return (1 - lam) * (a - b).abs().mean() + lam * (1 - m.mean())

    p = {k: torch.nn.Parameter(v) for k, v in train.raw_params_from_scene(sc, dev).items()}
    opt = torch.optim.Adam([{"params": [p[n]], "lr": lr, "name": n} for n, lr in train.DEFAULT_LRS], lr=0.0, eps=1e-15)
    max_radii2D, accum, denom = torch.zeros(P, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)
    gts_host = gt_images(min(nv, 4), H, W)
    gts = [t.to(dev) for t in gts_host]

    def iteration(cam, gt):
        means2D = torch.zeros_like(p["xyz"], requires_grad=True)
        image, radii, _, _ = RefRaster.apply(cam, p["xyz"], means2D, torch.cat((p["f_dc"], p["f_rest"]), dim=1),
                                             torch.sigmoid(p["opacity"]), torch.exp(p["scaling"]), F.normalize(p["rotation"]))
        loss = torch_loss(image.unsqueeze(0), gt.unsqueeze(0))
        loss.backward()
        vis = radii > 0
        max_radii2D[vis] = torch.max(max_radii2D[vis], radii[vis])
        accum[vis] += torch.norm(means2D.grad[vis, :2], dim=-1, keepdim=True)
        denom[vis] += 1
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss

    return torch, dev, iteration, cams, cd, gts_host, gts, (P, W, H, nv)


def run_reference_train(args, local, workload):
    torch, dev, iteration, cams, cd, gts_host, gts, (P, W, H, nv) = reference_trainer(local, workload)
    ms = _timed_single(torch, lambda i: iteration(cd[i % nv], gts[i % len(gts)]), args.steps, args.warmup)
    pre = InputPrefetcher(dev, (3, H, W), cams_to_host(cams), gts_host, lambda i: i % nv)
    res = ResultReader(torch)

    def e2e_step(i):
        v = i % nv
        sl = pre.take(i, False)
        cam = dict(cams[v], viewmatrix=sl["cam"][0:16], projmatrix=sl["cam"][16:32], campos=sl["cam"][32:35])
        loss = iteration(cam, sl["gt"])
        pre.release(sl)
        res.push(loss)

    pre.prefetch(0)
    ms_e2e = _timed_single(torch, e2e_step, args.steps, args.warmup, after=res.flush)
    return dict(P=P, W=W, H=H, nv=nv, ms_total=ms, ms_e2e=ms_e2e, h2d=3 * H * W * 4 + 35 * 4, d2h=4)


def run_reference_orbit(args, local, workload):
    torch, refmod, dev, sc, cams, g, cd, (P, W, H, nv) = _ref_inputs(workload, local)
    bg = torch.ones(3, device=dev)
    rr = refmod.RefRasterizer("dgr")
    outs = {}
    ms = _timed_single(torch, lambda i: rr.forward(g, cd[i % nv], bg, SH_DEGREE, outputs=outs), args.steps, args.warmup)
    pre = InputPrefetcher(dev, None, cams_to_host(cams), None, lambda i: i % nv)

    def e2e_step(i):   # utils/wheatgs_helper.py:290-291: render, then rendering.cpu() (synchronous)
        sl = pre.take(i, False)
        c = dict(cams[i % nv], viewmatrix=sl["cam"][0:16], projmatrix=sl["cam"][16:32], campos=sl["cam"][32:35])
        rr.forward(g, c, bg, SH_DEGREE, outputs=outs)
        pre.release(sl)
        outs["color"].cpu()

    pre.prefetch(0)
    ms_e2e = _timed_single(torch, e2e_step, args.steps, args.warmup)
    return dict(P=P, W=W, H=H, nv=nv, ms_total=ms, ms_e2e=ms_e2e, h2d=35 * 4, d2h=3 * H * W * 4)


def run_reference_votes(args, local, workload):
    """render_360.py:168-181 with the unmodified flashsplat rasterizer: a fresh zero-filled used_count[501, P] per
    view (FSR rasterize_points.cu:76), all_counts += used_count (kept on the GPU: kinder than the reference's .cpu()),
    then run_3d_seg.py:54-72 as the torch loop it is."""
    torch, refmod, dev, sc, cams, g, cd, (P, W, H, nv) = _ref_inputs(workload, local)
    nobj = WORKLOADS[workload]["labels"]
    bg = torch.from_numpy(BG).to(dev)
    masks_host = {v: torch.from_numpy(m).pin_memory() for v, m in make_masks(sc, cams, nobj, range(nv)).items()}
    masks = {v: m.to(dev) for v, m in masks_host.items()}
    rr = refmod.RefRasterizer("fsr")
    all_counts = torch.zeros((nobj + 1, P), device=dev)

    def step(i, mask=None, cam=None):
        v = i % nv
        o = rr.forward(g, cd[v] if cam is None else cam, bg, SH_DEGREE, gt_mask=masks[v] if mask is None else mask,
                       num_obj=nobj, outputs={})
        all_counts.add_(o["used_count"])

    ms = _timed_single(torch, step, args.steps, args.warmup)
    pre = InputPrefetcher(dev, (H, W), cams_to_host(cams), [masks_host[v] for v in range(nv)], lambda i: i % nv)

    def e2e_step(i):
        sl = pre.take(i, False)
        c = dict(cams[i % nv], viewmatrix=sl["cam"][0:16], projmatrix=sl["cam"][16:32], campos=sl["cam"][32:35])
        step(i, mask=sl["gt"], cam=c)
        pre.release(sl)

    pre.prefetch(0)
    ms_e2e = _timed_single(torch, e2e_step, args.steps, args.warmup)

    def ref_labels():
        s = all_counts.sum(dim=0)
        out = torch.zeros_like(all_counts).bool()
        for k in range(all_counts.shape[0]):
            oc = torch.stack([s - all_counts[k], all_counts[k]], dim=0)
            out[k] = torch.argmax(torch.nn.functional.normalize(oc, dim=0), dim=0)
        return out

    ms_labels = _timed_single(torch, lambda i: ref_labels(), 1, 1)
    return dict(P=P, W=W, H=H, nv=nv, ms_total=ms, ms_e2e=ms_e2e, ms_labels=ms_labels, h2d=H * W * 4 + 35 * 4, d2h=4)


def run_cpu_oracle(workload, budget_s=12.0, max_views=3):
    """CPU oracle (a port: oracle/splat_oracle.c, OpenMP over tiles / Gaussians) fwd+bwd on a bounded
    sample of the workload's views.  Returns (views/s, threads, sample text)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from wheat_3dgs_b200 import scene
    if WORKLOADS[workload]["kind"] != "fwdbwd" or workload == "cfg5":
        workload = "cfg2"   # the CPU port is timed on config 2's step whatever the GPU workload
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
METRICS = {
    "fwdbwd": ("train views/s (rasterizer fwd+bwd per view, %s; 1000/value = fwd+bwd ms/view)", "views/s"),
    "train": ("train iterations/s x views per iteration (full iteration: render, L1+SSIM loss, backward, statistics, "
              "Adam; %s)", "views/s"),
    "orbit": ("render MPix/s (orbit frames forward only with depth+alpha, %s)", "MPix/s"),
    "votes": ("flashsplat label-vote views/s (one ~500-id mask per view, votes accumulated for 501 x P, %s)", "views/s"),
}


def e2e_split(e2e_ms, device_ms, h2d):
    """What one end-to-end step is made of, so the line says what bounds it: the step as timed end to end, the same
    step with inputs resident (the line's ms_per_step), and the upload of one step's inputs as its own stream saw it
    (rank 0's; it runs beside the previous step's kernels).  The upload is prefetched one step ahead, so a step cannot
    be shorter than the longer of the two; what is left above that is host time between launches (Python, autograd).
    bound: "h2d" / "device" = the longer of the two legs, when the step is within 10 % of it; else "host"."""
    out = {"e2e_ms_per_step": e2e_ms, "device_ms_per_step": device_ms}
    if not h2d:
        return out
    out.update(h2d_ms_per_step=h2d["h2d_ms_mean"], h2d_ms_max=h2d["h2d_ms_max"], h2d_gbs=h2d["h2d_gbs"])
    floor = max(h2d["h2d_ms_mean"], device_ms)
    out["bound"] = "host" if e2e_ms > 1.10 * floor else ("h2d" if h2d["h2d_ms_mean"] >= device_ms else "device")
    out["over_floor_ms"] = e2e_ms - floor
    return out


def describe(workload):
    wl = WORKLOADS[workload]
    shape = "%dK Gaussians SH3, %dx%d" % (wl["P"] // 1000, wl["W"], wl["H"]) if wl["P"] < 1000000 else \
        "%dM Gaussians SH3, %dx%d" % (wl["P"] // 1000000, wl["W"], wl["H"])
    what = {"fwdbwd": "%d overhead views, fwd+bwd per view", "train": "%d overhead views, full training iteration per view",
            "orbit": "%d orbit frames (elevation 45, distance 2 x scene radius), forward only",
            "votes": "%d overhead views x one ~500-id mask, flashsplat forward per view"}[wl["kind"]] % wl["views"]
    return shape, "%s: synthetic plot_461-shaped scene, P=%d Gaussians SH3, %dx%d, %s" % (workload, wl["P"], wl["W"], wl["H"], what)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=144)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cpu"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="cfg2 only: skip the train / orbit legs of `extra`")
    ap.add_argument("--seeds", type=int, default=6, help="cfg3: seed masks per rank for the run_3d_seg loop")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]
    kind = wl["kind"]
    shape, wtext = describe(args.workload)
    metric, unit = METRICS[kind][0] % shape, METRICS[kind][1]
    # identical in both arms (the driver compares the dicts); how the N ranks are used is a separate key
    config = {"workload": wtext, "units_per_step_per_gpu": 1,
              "l2": "per-step inputs (Gaussians + instance lists, > 800 MB at 1M Gaussians) exceed the 126 MB L2; no flush"}
    base = {"metric": metric, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config}

    if args.impl != "ours":
        if rank != 0:
            return 0
        # the reference is single-GPU (utils/general_utils.py:133): under torchrun rank 0 alone runs it; n_gpus is the
        # launch's N so that the line pairs with this arm's line of the same N
        line = dict(base, impl="reference", n_gpus=world,
                    parallelism="reference is single-GPU: one unit per step on rank 0"
                    + (" (launched with %d ranks, the others idle)" % world if world > 1 else ""))
        if args.impl == "reference":
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import ref as refmod
            need = "fsr" if kind == "votes" else "dgr"
            if not refmod.available(need):
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_%s.so not built "
                                  "(python oracle/build_ref.py needs /root/reference)" % need}))
                return 0
            if kind == "fwdbwd":
                r = run_reference_fwdbwd(args, local)
            elif kind == "train":
                r = run_reference_train(args, local, args.workload)
            elif kind == "orbit":
                r = run_reference_orbit(args, local, args.workload)
            else:
                r = run_reference_votes(args, local, args.workload)
            scale = (r["W"] * r["H"] / 1e6) if kind == "orbit" else 1.0
            v = scale * args.steps / (r["ms_total"] * 1e-3)
            line.update(value=v, ms_per_step=r["ms_total"] / args.steps, clocks=r.get("clocks"), gpu_launches=None,
                        reference_kind="unmodified reference CUDA rasterizer (%s) built for sm_100, same B200; the "
                        "reference has no CPU path" % ("flashsplat-rasterization" if kind == "votes" else
                                                       "diff-gaussian-rasterization @8829d14f"),
                        e2e={"value": scale * args.steps / (r["ms_e2e"] * 1e-3), "unit": unit,
                             "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"]})
            if "ms_labels" in r:
                line["extra"] = {"multi_instance_opt_ms": r["ms_labels"]}
            if not args.no_cpu_baseline and kind == "fwdbwd":
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
    job = Job(rank, world, local)
    par = {"fwdbwd": "views sharded x%d (views of similar instance count share a step), Gaussians replicated, 1 in-place "
                     "SUM of 62*P fp32 per step" % world,
           "train": "one view per rank per iteration, Gaussians replicated, gradients reduced + Adam + parameters "
                    "broadcast by one kernel per iteration",
           "orbit": "frames rank::%d, no exchange" % world,
           "votes": "views rank::%d, ONE sum of used_count[501,P] over ranks at the end of the pass" % world}[kind] \
        if world > 1 else "single GPU"

    if kind == "fwdbwd":
        r = run_fwdbwd(args, job)
        extra = {}
        if args.workload == "cfg2" and not args.no_extra:
            # BASELINE.json's other two metric parts, same run, same N, reduced step counts
            tr = run_train(args, job, "train", min(args.steps, 36), args.warmup, with_e2e=False)
            ob = run_orbit(args, job, "cfg4", min(args.steps, 40), args.warmup, with_e2e=False)
            extra = {"train_iters_s": 1000.0 * min(args.steps, 36) / tr["ms_total"],
                     "train_views_s": world * 1000.0 * min(args.steps, 36) / tr["ms_total"],
                     "train_how": tr["how"], "train_parity": tr["parity"],
                     "render_mpix_s": world * min(args.steps, 40) * ob["W"] * ob["H"] / 1e6 / (ob["ms_total"] * 1e-3),
                     "render_ms_per_frame": ob["ms_total"] / min(args.steps, 40), "render_parity": ob["parity"]}
        if rank != 0:
            if world > 1:
                torch.distributed.destroy_process_group()
            return 0
        P, W, H = r["P"], r["W"], r["H"]
        peaks = load_peaks()
        st = r["stats"]
        T = ((W + 15) // 16) * ((H + 15) // 16)
        bits = max(1, int(np.ceil(np.log2(T + 1))))  # tile-id bits sorted per instance (depth-first binning)
        by, fl = stage_model(P, st["V"], st["R"], T, W * H, r["M"], (bits + 7) // 8, st["n_eval"], st["n_contrib"],
                             Rf=st["R"] - st["instances_never_fetched"])
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
        traffic, traffic_src = None, None  # DRAM bytes per launch of the dominant kernel, from a committed ncu --set full capture
        for tf in ("r02_traffic.json", "r01_traffic.json"):
            try:
                with open(os.path.join(ROOT, "profiles", tf)) as f:
                    tj = json.load(f)
                if args.workload == "cfg2" and top in tj["bytes_per_launch"]:
                    traffic = tj["bytes_per_launch"][top]
                    traffic_src = "profiles/%s: %s" % (tf, tj.get("note", "ncu --set full, config 2 view 17, one launch"))
                    break
            except (OSError, ValueError, KeyError):
                pass
        sm_mhz = r["clocks"].get("sm_mhz") or peaks["sm_max_mhz"]
        if top in fl:
            nominal = 148 * 128 * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12
            peak = r["fma"]["ffma2_tflops"]
            roof = {"kernel": top, "bound": "fp32", "achieved": stages[top]["tflops"], "peak": peak, "unit": "TFLOP/s",
                    "frac": stages[top]["tflops"] / peak, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": "measured in this run: packed FFMA2 chains, 3 register operands (ws_debug_fma_peak); "
                                   "scalar FFMA %.1f TFLOP/s; nominal 148 SM x 128 lanes x 2 x %.0f MHz = %.1f"
                                   % (r["fma"]["ffma_tflops"], peaks["sm_max_mhz"], nominal),
                    "frac_of_nominal": stages[top]["tflops"] / nominal,
                    "frac_at_observed_clock": stages[top]["tflops"] / (148 * 128 * 2 * sm_mhz * 1e6 / 1e12),
                    "hbm_gbs": stages[top]["gbs"], "hbm_frac": stages[top]["hbm_frac"], "issue_active": issue_active(top)}
        else:
            roof = {"kernel": top, "bound": "hbm", "achieved": stages[top]["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": stages[top]["hbm_frac"], "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": peaks["source"]}
        value = world * args.steps / (r["ms_total"] * 1e-3)
        line = dict(base, impl="ours", value=value, ms_per_step=r["ms_total"] / args.steps, clocks=r["clocks"],
                    gpu_launches=r["launches"], roofline=roof, stages=stages, parallelism=par,
                    work_per_view=dict({k: st[k] for k in ("V", "R", "n_eval", "n_contrib", "quad_pairs",
                                                           "instances_never_fetched")},
                                       quad_cull_survival=st["quad_pairs"] / max(4.0 * st["R"], 1.0),
                                       pixel_slots_per_contributing_pair=64.0 * st["quad_pairs"] / max(st["n_contrib"], 1.0)),
                    parity=r["parity"],
                    e2e={"value": world * args.steps / (r["ms_e2e"] * 1e-3), "unit": "views/s",
                         "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"]})
        line["e2e"].update(e2e_split(r["ms_e2e"] / args.steps, r["ms_total"] / args.steps, r.get("h2d_stats")))
        if extra:
            line["extra"] = extra
        if world > 1:
            line["all_reduce"] = r["reduce_how"]
            line["exchange_check"] = r["exchange_check"]
            line["exchange_split"] = r["exchange_split"]
        if world == 1 and not args.no_cpu_baseline:
            cv, cores, sample = run_cpu_oracle(args.workload)
            line["cpu_baseline"] = {"value": cv, "unit": "views/s", "cores": cores, "kind": "port", "sample": sample}
    else:
        clk = ClockSampler(job.local)
        if rank == 0:
            clk.start()
            clk.wait_started()
        t0 = time.time()
        if kind == "train":
            r = run_train(args, job, args.workload, args.steps, args.warmup)
            scale = 1.0
        elif kind == "orbit":
            r = run_orbit(args, job, args.workload, args.steps, args.warmup)
            scale = r["W"] * r["H"] / 1e6
        else:
            r = run_votes(args, job, args.workload, args.steps, args.warmup)
            scale = 1.0
        clocks = clk.stop(t0, time.time())
        if rank != 0:
            if world > 1:
                torch.distributed.destroy_process_group()
            return 0
        value = scale * world * args.steps / (r["ms_total"] * 1e-3)
        line = dict(base, impl="ours", value=value, ms_per_step=r["ms_total"] / args.steps, clocks=clocks,
                    gpu_launches=r["launches"], parallelism=par, parity=r["parity"],
                    e2e={"value": scale * world * args.steps / (r["ms_e2e"] * 1e-3), "unit": unit,
                         "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"]})
        if kind == "train":
            line["extra"] = {"train_iters_s": 1000.0 * args.steps / r["ms_total"], "exchange_and_step": r["how"],
                             "final_loss": r["final_loss"]}
        elif kind == "orbit":
            line["extra"] = {"ms_per_frame": r["ms_total"] / args.steps}
            if "ms_gathered" in r:
                line["extra"]["gathered_mpix_s"] = scale * world * args.steps / (r["ms_gathered"] * 1e-3)
        else:
            line["extra"] = {"vote_exchange": r["how"], "vote_exchange_ms": r["ms_reduce"], "vote_rows": r["rows"],
                             "multi_instance_opt_ms": r["ms_labels"], "labelled_gaussians": r["labelled"],
                             "run_3d_seg": r["seg"]}
    print(json.dumps(line))
    if world > 1:
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
