"""One full training iteration of train_vanilla_3dgs.py:65-112 on config 2 (1 M Gaussians SH3, 2048x1536,
36 views), timed with CUDA events, two stacks on the same GPU and the same raw parameters:

  ours-raw   as ours, but the getters (exp / sigmoid / normalize / cat) run inside preprocess (forward_raw)
  ours       torch activations -> drop-in GaussianRasterizer (libwheatsplat) -> fused L1+SSIM loss (ws_loss.cu)
             -> backward -> fused densification statistics + fused Adam (ws_optim.cu)
  reference  torch activations -> UNMODIFIED reference rasterizer (oracle/_ref/libref_dgr.so behind an autograd
             Function with the reference's zero fills) -> utils/loss_utils.py semantics in torch (5 grouped
             conv2d) -> backward -> the reference's masked-index statistics + torch.optim.Adam

    python tools/train_iter.py [iters]        # prints iterations/s of both and the per-part split of ours
    torchrun --nproc-per-node N ... tools/train_iter.py [iters] ours-raw
                                              # data parallel: every rank its own view per iteration,
                                              # gradients written into the symmetric bucket by the backward, summed by
                                              # libwheatsplat's exchange kernel, then the fused Adam on every rank
"""
import math
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import wheat_3dgs_b200  # noqa: E402
from wheat_3dgs_b200 import dist as wdist, loss as wloss, optim as woptim, rasterizer as wrast, scene  # noqa: E402

P, W, H, DEG = 1000000, 2048, 1536, 3
ITERS = int(sys.argv[1]) if len(sys.argv) > 1 else 36
WORLD = int(os.environ.get("WORLD_SIZE", "1"))
RANK = int(os.environ.get("RANK", "0"))
if WORLD > 1:
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    wdist.init_from_env()
dev = torch.device("cuda", torch.cuda.current_device())
sc = scene.synthetic_scene(P, seed=0)
cams = scene.overhead_cameras(W, H)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
cd = [dict(c, viewmatrix=t(c["viewmatrix"]), projmatrix=t(c["projmatrix"]), campos=t(c["campos"])) for c in cams]
bg = torch.zeros(3, device=dev)
gts = [torch.rand(3, H, W, device=dev) for _ in range(4)]
LRS = [("xyz", 0.00016), ("f_dc", 0.0025), ("f_rest", 0.0025 / 20), ("opacity", 0.05), ("scaling", 0.005),
       ("rotation", 0.001)]


def raw_params():
    """GaussianModel's raw parameters (scene/gaussian_model.py:147-161): log scales, opacity logits, raw rotations."""
    op = np.clip(sc["opacities"], 1e-6, 1 - 1e-6)
    raw = {"xyz": sc["means3D"], "f_dc": sc["shs"][:, :1, :], "f_rest": sc["shs"][:, 1:, :],
           "opacity": np.log(op / (1 - op)), "scaling": np.log(sc["scales"]), "rotation": sc["rotations"] * 1.7}
    return {k: torch.nn.Parameter(t(v.astype(np.float32))) for k, v in raw.items()}


def activations(p):
    """GaussianModel getters (scene/gaussian_model.py:101-121)."""
    return dict(means3D=p["xyz"], shs=torch.cat((p["f_dc"], p["f_rest"]), dim=1), opacities=torch.sigmoid(p["opacity"]),
                scales=torch.exp(p["scaling"]), rotations=F.normalize(p["rotation"]))


# ---------------------------------------------------------------- reference stack
class _RefRaster(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rr, cam, means3D, means2D, shs, opacities, scales, rotations):
        g = dict(means3D=means3D.detach(), shs=shs.detach(), opacities=opacities.detach(), scales=scales.detach(),
                 rotations=rotations.detach())
        o = rr.forward(g, cam, bg, DEG, outputs={})
        ctx.rr = rr
        ctx.mark_non_differentiable(o["radii"])
        return o["color"], o["radii"], o["depth"], o["alpha"]

    @staticmethod
    def backward(ctx, gc, gr, gd, ga):
        z = lambda x, ref: torch.zeros_like(ref) if x is None else x.contiguous()
        o = ctx.rr.last["out"]
        g = ctx.rr.backward(z(gc, o["color"]), z(gd, o["depth"]), z(ga, o["alpha"]), grads={})
        return (None, None, g["dL_dmeans3D"], g["dL_dmeans2D"], g["dL_dsh"], g["dL_dopacity"], g["dL_dscales"],
                g["dL_drotations"])


g1 = torch.tensor([math.exp(-(i - 5) ** 2 / 4.5) for i in range(11)])
g1 = (g1 / g1.sum()).unsqueeze(1)
WIN = g1.mm(g1.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().to(dev)


def torch_loss(a, b, lam=0.2):
    """utils/loss_utils.py:17-62 as train_vanilla_3dgs.py:77-79 combines them."""
    mu1, mu2 = F.conv2d(a, WIN, padding=5, groups=3), F.conv2d(b, WIN, padding=5, groups=3)
    s1 = F.conv2d(a * a, WIN, padding=5, groups=3) - mu1 * mu1
    s2 = F.conv2d(b * b, WIN, padding=5, groups=3) - mu2 * mu2
    s12 = F.conv2d(a * b, WIN, padding=5, groups=3) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    return (1 - lam) * (a - b).abs().mean() + lam * (1 - m.mean())


class Trainer:
    def __init__(self, kind):
        self.kind = kind
        self.p = raw_params()
        groups = [{"params": [self.p[n]], "lr": lr, "name": n} for n, lr in LRS]
        self.opt = (woptim.Adam if kind.startswith("ours") else torch.optim.Adam)(groups, lr=0.0, eps=1e-15)
        self.max_radii2D = torch.zeros(P, device=dev)
        self.accum = torch.zeros(P, 1, device=dev)
        self.denom = torch.zeros(P, 1, device=dev)
        if kind.startswith("ours"):
            wheat_3dgs_b200.install()
            from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
            self.RS, self.R = GaussianRasterizationSettings, GaussianRasterizer
        else:
            import ref as refmod
            self.rr = refmod.RefRasterizer("dgr")
        self.parts = {}
        self.reduce = None
        self.sharded = None
        if WORLD > 1:
            if kind not in ("ours-raw", "ours-fused"):
                raise SystemExit("data-parallel mode runs the ours-raw / ours-fused stacks")
            self.n_train = wdist.train_bucket_numel(P, 16)
            if kind == "ours-fused":
                # exchange + Adam as ONE kernel (dist.ShardedAdam): parameters in symmetric memory, rank r updates slice r
                sb = wdist.SymmetricBucket(wdist.full_bucket_numel(P, 16), dev)
                offs = {"xyz": 0, "f_dc": 6 * P, "f_rest": 9 * P, "opacity": 54 * P, "scaling": 55 * P, "rotation": 58 * P}
                self.sharded = wdist.ShardedAdam([{"name": n, "tensor": self.p[n], "lr": lr, "grad_offset": offs[n]}
                                                  for n, lr in LRS], sb, eps=1e-15)
                self.p = self.sharded.params
                self.bucket, self.reduce = sb.tensor, None
                self.how = "fused reduce+Adam+broadcast, " + ("multimem" if self.sharded.param_mc else "peers")
            else:
                self.bucket, self.reduce, self.how = wdist.make_bucket(wdist.full_bucket_numel(P, 16), dev)
            wrast.set_grad_bucket(self.bucket)

    def iteration(self, i, ev=None):
        mark = (lambda name: ev.append((name, self._rec()))) if ev is not None else (lambda name: None)
        c = cd[(RANK + i * WORLD) % len(cd)]
        gt = gts[i % len(gts)]
        if self.kind in ("ours-raw", "ours-fused"):
            means2D = torch.zeros_like(self.p["xyz"], requires_grad=True)
            mark("activations")
            rs = self.RS(image_height=H, image_width=W, tanfovx=c["tanfovx"], tanfovy=c["tanfovy"], bg=bg,
                         scale_modifier=1.0, viewmatrix=c["viewmatrix"], projmatrix=c["projmatrix"], sh_degree=DEG,
                         campos=c["campos"], prefiltered=False, debug=False)
            p = self.p
            image, radii, depth, alpha = self.R(rs).forward_raw(p["xyz"], means2D, p["f_dc"], p["f_rest"], p["opacity"],
                                                                p["scaling"], p["rotation"])
            mark("render_fwd")
            loss = wloss.l1_ssim_loss(image, gt, 0.2)
            mark("loss_fwd")
            loss.backward()
            mark("backward")
            woptim.add_densification_stats(radii, means2D.grad, self.max_radii2D, self.accum, self.denom)
            mark("densify_stats")
            if self.reduce is not None:
                # the statistics above are this rank's (summed over ranks only when densification reads them);
                # the parameter gradients are the bucket's slices, summed in place over all ranks
                self.reduce(self.n_train)
                b = self.bucket
                p["xyz"].grad, p["f_dc"].grad = b[0:3 * P].view(P, 3), b[6 * P:9 * P].view(P, 1, 3)
                p["f_rest"].grad, p["opacity"].grad = b[9 * P:54 * P].view(P, 15, 3), b[54 * P:55 * P].view(P, 1)
                p["scaling"].grad, p["rotation"].grad = b[55 * P:58 * P].view(P, 3), b[58 * P:62 * P].view(P, 4)
                mark("exchange")
            if self.sharded is not None:
                self.sharded.step()
                for q in p.values():
                    q.grad = None
                mark("exchange+adam")
                return loss
            self.opt.step()
            self.opt.zero_grad(set_to_none=True)
            mark("adam")
            return loss
        a = activations(self.p)
        means2D = torch.zeros_like(a["means3D"], requires_grad=True)
        mark("activations")
        if self.kind == "ours":
            rs = self.RS(image_height=H, image_width=W, tanfovx=c["tanfovx"], tanfovy=c["tanfovy"], bg=bg,
                         scale_modifier=1.0, viewmatrix=c["viewmatrix"], projmatrix=c["projmatrix"], sh_degree=DEG,
                         campos=c["campos"], prefiltered=False, debug=False)
            image, radii, depth, alpha = self.R(rs)(means3D=a["means3D"], means2D=means2D, opacities=a["opacities"],
                                                   shs=a["shs"], scales=a["scales"], rotations=a["rotations"])
            mark("render_fwd")
            loss = wloss.l1_ssim_loss(image, gt, 0.2)
        else:
            image, radii, depth, alpha = _RefRaster.apply(self.rr, c, a["means3D"], means2D, a["shs"], a["opacities"],
                                                          a["scales"], a["rotations"])
            mark("render_fwd")
            loss = torch_loss(image.unsqueeze(0), gt.unsqueeze(0))
        mark("loss_fwd")
        loss.backward()
        mark("backward")
        if self.kind == "ours":
            woptim.add_densification_stats(radii, means2D.grad, self.max_radii2D, self.accum, self.denom)
        else:
            vis = radii > 0
            self.max_radii2D[vis] = torch.max(self.max_radii2D[vis], radii[vis])
            self.accum[vis] += torch.norm(means2D.grad[vis, :2], dim=-1, keepdim=True)
            self.denom[vis] += 1
        mark("densify_stats")
        self.opt.step()
        self.opt.zero_grad(set_to_none=True)
        mark("adam")
        return loss

    @staticmethod
    def _rec():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e


def sync():
    if WORLD > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()


def run(kind):
    tr = Trainer(kind)
    for i in range(3):
        tr.iteration(i)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(ITERS):
        loss = tr.iteration(3 + i)
    e1.record()
    sync()
    ms = e0.elapsed_time(e1) / ITERS
    if WORLD > 1:
        ms = wdist.max_over_ranks(ms, dev)
    # per-part split (second pass, events between the parts)
    parts = {}
    for i in range(8):
        ev = [("start", Trainer._rec())]
        tr.iteration(i, ev)
        torch.cuda.synchronize()
        for (n0, a), (n1, b) in zip(ev[:-1], ev[1:]):
            parts[n1] = parts.get(n1, 0.0) + a.elapsed_time(b) / 8
    if RANK == 0:
        if WORLD > 1:
            print("%-9s %d GPUs (%s): %.3f ms/iteration = %.1f iterations/s = %.1f views/s   loss %.5f"
                  % (kind, WORLD, tr.how, ms, 1000.0 / ms, WORLD * 1000.0 / ms, float(loss)))
        else:
            print("%-9s %.3f ms/iteration = %.1f iterations/s   loss %.5f" % (kind, ms, 1000.0 / ms, float(loss)))
        print("          " + "  ".join("%s %.3f" % kv for kv in parts.items()))
    wrast.set_grad_bucket(None)
    del tr
    torch.cuda.empty_cache()
    return ms


if __name__ == "__main__":
    if len(sys.argv) > 2:            # one stack only (profiling target): python tools/train_iter.py 3 ours-raw
        run(sys.argv[2])
        if WORLD > 1:
            torch.distributed.destroy_process_group()
        sys.exit(0)
    a = run("ours")
    r = run("ours-raw")
    b = run("reference")
    print("full-iteration speed-up x%.2f (drop-in API), x%.2f (raw-parameter entry)" % (b / a, b / r))
