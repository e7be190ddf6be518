"""Host mirror of the training iteration around the rasterizer (train_vanilla_3dgs.py:55-115), on this library's
fused stages:

    render            GaussianRasterizer.forward_raw: the getters' activations (scene/gaussian_model.py:101-121)
                      run inside preprocess                                        train_vanilla_3dgs.py:73
    loss              0.8 L1 + 0.2 (1 - SSIM), one fused forward + backward        :77-80 (utils/loss_utils.py:17-62)
    statistics        max_radii2D / xyz_gradient_accum / denom in one pass         :100-103
    exchange + step   N GPUs: every rank renders its own view, the gradients land in a symmetric-memory bucket and
                      ONE kernel reduces rank r's slice through the switch, applies Adam to it and stores the new
                      parameters into every replica (dist.ShardedAdam); 1 GPU: the fused Adam (optim.Adam)   :110-112

`Trainer.iteration(cam, gt)` is one iteration without the every-100-iterations densification
(densify.densify_and_prune / reset_opacity, called by the caller as the reference does at :105-109).
No CPU path: everything lives on the rank's GPU.
"""
import numpy as np
import torch

from . import dist as wdist
from . import loss as wloss
from . import optim as woptim
from . import rasterizer as wrast

# arguments/__init__.py:74-82 defaults (position_lr_init is rescheduled every iteration by update_learning_rate)
DEFAULT_LRS = (("xyz", 0.00016), ("f_dc", 0.0025), ("f_rest", 0.0025 / 20), ("opacity", 0.05), ("scaling", 0.005),
               ("rotation", 0.001))


def raw_params_from_scene(sc, dev):
    """GaussianModel's raw parameters (scene/gaussian_model.py:147-161) from a post-activation synthetic scene:
    log scales, opacity logits, un-normalised rotations."""
    op = np.clip(sc["opacities"], 1e-6, 1 - 1e-6)
    raw = {"xyz": sc["means3D"], "f_dc": sc["shs"][:, :1, :], "f_rest": sc["shs"][:, 1:, :],
           "opacity": np.log(op / (1 - op)), "scaling": np.log(sc["scales"]), "rotation": sc["rotations"] * 1.7}
    return {k: torch.from_numpy(np.ascontiguousarray(v.astype(np.float32))).to(dev) for k, v in raw.items()}


class Trainer:
    def __init__(self, raw, sh_degree, device, lrs=DEFAULT_LRS, lambda_dssim=0.2, world=1, fused_exchange=True):
        self.dev, self.deg, self.lam, self.world = device, int(sh_degree), float(lambda_dssim), int(world)
        self.P = int(raw["xyz"].shape[0])
        P = self.P
        self.M = 1 + int(raw["f_rest"].shape[1])
        self.p = {k: torch.nn.Parameter(v.detach().clone().to(device)) for k, v in raw.items()}
        self.max_radii2D = torch.zeros(P, device=device)
        self.xyz_gradient_accum = torch.zeros(P, 1, device=device)
        self.denom = torch.zeros(P, 1, device=device)
        self.sharded, self.reduce, self.bucket, self.opt = None, None, None, None
        self.how = "single GPU, fused Adam"
        M = self.M
        # offsets of the six parameter gradients inside the flat bucket ops._backward() fills
        self._off = {"xyz": 0, "f_dc": 6 * P, "f_rest": 9 * P, "opacity": (6 + 3 * M) * P, "scaling": (7 + 3 * M) * P,
                     "rotation": (10 + 3 * M) * P}
        self._shape = {k: tuple(v.shape) for k, v in self.p.items()}
        if self.world > 1:
            self.n_train = wdist.train_bucket_numel(P, M)
            if fused_exchange and P % 4 == 0:
                sb = wdist.SymmetricBucket(wdist.full_bucket_numel(P, M), device)
                self.sharded = wdist.ShardedAdam([{"name": n, "tensor": self.p[n], "lr": lr, "grad_offset": self._off[n]}
                                                  for n, lr in lrs], sb, eps=1e-15)
                self.p = self.sharded.params
                self.bucket = sb.tensor
                self.how = "reduce-scatter + Adam + all-gather in one kernel (%s)" % (
                    "multimem" if self.sharded.param_mc else "peer loads/stores")
            else:
                self.bucket, self.reduce, how = wdist.make_bucket(wdist.full_bucket_numel(P, M), device)
                self.how = how + " + fused Adam on every rank"
        if self.sharded is None:
            groups = [{"params": [self.p[n]], "lr": lr, "name": n} for n, lr in lrs]
            self.opt = woptim.Adam(groups, lr=0.0, eps=1e-15)   # scene/gaussian_model.py:181
        from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer
        self._RS, self._R = GaussianRasterizationSettings, GaussianRasterizer

    def render(self, cam, bg, means2D):
        rs = self._RS(image_height=cam["image_height"], image_width=cam["image_width"], tanfovx=cam["tanfovx"],
                      tanfovy=cam["tanfovy"], bg=bg, scale_modifier=1.0, viewmatrix=cam["viewmatrix"],
                      projmatrix=cam["projmatrix"], sh_degree=self.deg, campos=cam["campos"], prefiltered=False,
                      debug=False)
        p = self.p
        return self._R(rs).forward_raw(p["xyz"], means2D, p["f_dc"], p["f_rest"], p["opacity"], p["scaling"],
                                       p["rotation"])

    def iteration(self, cam, gt, bg):
        """One training iteration on this rank's view; returns the loss (device scalar)."""
        p = self.p
        means2D = torch.zeros_like(p["xyz"], requires_grad=True)
        if self.bucket is not None:
            wrast.set_grad_bucket(self.bucket)
        try:
            image, radii, _, _ = self.render(cam, bg, means2D)
            loss = wloss.l1_ssim_loss(image, gt, self.lam)
            loss.backward()
        finally:
            if self.bucket is not None:
                wrast.set_grad_bucket(None)
        woptim.add_densification_stats(radii, means2D.grad, self.max_radii2D, self.xyz_gradient_accum, self.denom)
        if self.sharded is not None:
            self.sharded.step()
            for q in p.values():
                q.grad = None
            return loss.detach()
        if self.reduce is not None:
            # the statistics above stay per rank (summed only when densification reads them,
            # dist.allreduce_densify_stats); the parameter gradients are the bucket's slices, summed in place
            self.reduce(self.n_train)
            b = self.bucket
            for n in p:
                numel = p[n].numel()
                p[n].grad = b[self._off[n]:self._off[n] + numel].view(self._shape[n])
        self.opt.step()
        self.opt.zero_grad(set_to_none=True)
        return loss.detach()
