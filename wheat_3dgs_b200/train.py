"""Host mirror of the training iteration around the rasterizer (train_vanilla_3dgs.py:55-115), on this library's
fused stages:

    render            GaussianRasterizer.forward_raw: the getters' activations (scene/gaussian_model.py:101-121)
                      run inside preprocess                                        train_vanilla_3dgs.py:73
    loss              0.8 L1 + 0.2 (1 - SSIM), one fused forward + backward        :77-80 (utils/loss_utils.py:17-62)
    statistics        max_radii2D / xyz_gradient_accum / denom in one pass         :100-103
    exchange + step   N GPUs: every rank renders its own view, the gradients land in a symmetric-memory bucket and
                      ONE kernel reduces rank r's slice through the switch, applies Adam to it and stores the new
                      parameters into every replica (dist.ShardedAdam); 1 GPU: the fused Adam (optim.Adam)   :110-112

`Trainer.iteration(cam, gt, bg)` is one iteration without the every-100-iterations densification
(densify.densify_and_prune / reset_opacity, called by the caller as the reference does at :105-109).
Pass `iteration=i` (1-based, as the reference counts) and the top of the loop body runs too: the position learning
rate of that iteration (update_learning_rate, :58) and one more SH band every 1000 iterations (oneupSHdegree, :61-62).
Without it the caller owns both (`update_learning_rate`, `oneup_sh_degree`, `set_lr`).
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


# arguments/__init__.py:74-77
POSITION_LR = dict(lr_init=0.00016, lr_final=0.0000016, lr_delay_mult=0.01, max_steps=30_000)


def expon_lr(step, lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """The learning-rate curve of utils/general_utils.py:29-62 at `step`: log-linear from lr_init (step 0) to lr_final
    (step >= max_steps), times a sine ease-in from lr_delay_mult over the first lr_delay_steps steps when that is > 0.
    Same float64 operations in the same order as the reference, so the rates are equal to the last bit."""
    if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
        return 0.0
    ease = 1.0
    if lr_delay_steps > 0:
        ease = lr_delay_mult + (1 - lr_delay_mult) * np.sin(0.5 * np.pi * np.clip(step / lr_delay_steps, 0, 1))
    t = np.clip(step / max_steps, 0, 1)
    return float(ease * np.exp(np.log(lr_init) * (1 - t) + np.log(lr_final) * t))


def raw_params_from_scene(sc, dev):
    """GaussianModel's raw parameters (scene/gaussian_model.py:147-161) from a post-activation synthetic scene:
    log scales, opacity logits, un-normalised rotations."""
    op = np.clip(sc["opacities"], 1e-6, 1 - 1e-6)
    raw = {"xyz": sc["means3D"], "f_dc": sc["shs"][:, :1, :], "f_rest": sc["shs"][:, 1:, :],
           "opacity": np.log(op / (1 - op)), "scaling": np.log(sc["scales"]), "rotation": sc["rotations"] * 1.7}
    return {k: torch.from_numpy(np.ascontiguousarray(v.astype(np.float32))).to(dev) for k, v in raw.items()}


class Trainer:
    def __init__(self, raw, sh_degree, device, lrs=DEFAULT_LRS, lambda_dssim=0.2, world=1, fused_exchange=True,
                 spatial_lr_scale=1.0, position_lr=POSITION_LR):
        """`sh_degree`: the ACTIVE degree the run starts with (the reference starts at 0 and raises it every 1000
        iterations up to the degree the coefficient count allows; the benchmarks start at 3).  `spatial_lr_scale`:
        the scene extent factor the reference multiplies into the position rate (scene/gaussian_model.py:173,182-183);
        the "xyz" entry of `lrs` is taken as already scaled, the schedule in `position_lr` is scaled here."""
        self.dev, self.deg, self.lam, self.world = device, int(sh_degree), float(lambda_dssim), int(world)
        self.P = int(raw["xyz"].shape[0])
        P = self.P
        self.M = 1 + int(raw["f_rest"].shape[1])
        self.max_sh_degree = int(round(self.M ** 0.5)) - 1
        self._xyz_lr = dict(position_lr, lr_init=position_lr["lr_init"] * spatial_lr_scale,
                            lr_final=position_lr["lr_final"] * spatial_lr_scale)
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

    @property
    def active_sh_degree(self):
        return self.deg

    def oneup_sh_degree(self):
        """GaussianModel.oneupSHdegree (scene/gaussian_model.py:134-136)."""
        if self.deg < self.max_sh_degree:
            self.deg += 1

    def set_lr(self, name, lr):
        """param_group["lr"] of the group called `name`, whichever optimiser path this trainer runs."""
        if self.sharded is not None:
            self.sharded.set_lr(name, lr)
            return
        for g in self.opt.param_groups:
            if g["name"] == name:
                g["lr"] = float(lr)

    def get_lr(self, name):
        groups = self.sharded.groups if self.sharded is not None else self.opt.param_groups
        for g in groups:
            if g["name"] == name:
                return g["lr"]
        raise KeyError(name)

    def update_learning_rate(self, iteration):
        """GaussianModel.update_learning_rate (scene/gaussian_model.py:188-194): the position rate of `iteration`."""
        lr = expon_lr(iteration, **self._xyz_lr)
        self.set_lr("xyz", lr)
        return lr

    def render(self, cam, bg, means2D):
        rs = self._RS(image_height=cam["image_height"], image_width=cam["image_width"], tanfovx=cam["tanfovx"],
                      tanfovy=cam["tanfovy"], bg=bg, scale_modifier=1.0, viewmatrix=cam["viewmatrix"],
                      projmatrix=cam["projmatrix"], sh_degree=self.deg, campos=cam["campos"], prefiltered=False,
                      debug=False)
        p = self.p
        return self._R(rs).forward_raw(p["xyz"], means2D, p["f_dc"], p["f_rest"], p["opacity"], p["scaling"],
                                       p["rotation"])

    def iteration(self, cam, gt, bg, iteration=None):
        """One training iteration on this rank's view; returns the loss (device scalar).  `iteration` (1-based): also
        do what the reference does at the top of its loop body (train_vanilla_3dgs.py:58-62)."""
        if iteration is not None:
            self.update_learning_rate(iteration)
            if iteration % 1000 == 0:
                self.oneup_sh_degree()
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
