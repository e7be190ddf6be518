"""GPU: the fused Adam step and densification statistics (csrc/ws_optim.cu through wheat_3dgs_b200.optim, C ABI
ws_adam_step / ws_densify_stats) against torch.optim.Adam + the reference's own add_densification_stats (goldens
made by tests/golden/make_golden_optim.py), against the CPU oracle, and against torch.optim.Adam's CUDA
multi-tensor path at the training size."""
import os

import numpy as np
import pytest
import torch

import cases
import optim_oracle

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from wheat_3dgs_b200 import optim as woptim


def _close(name, a, b):
    """Optimiser tier (float32, ~1 ulp per operation; exp_avg cancels, hence the absolute floor)."""
    a, b = np.asarray(a, np.float32).reshape(-1), np.asarray(b, np.float32).reshape(-1)
    lim = 2e-6 * np.abs(b) + 2e-7 * np.abs(b).max()
    bad = np.abs(a - b) > lim
    assert not bad.any(), "%s: %d elements off, max err %g" % (name, bad.sum(), np.abs(a - b).max())


def _groups(params):
    tp = {n: torch.nn.Parameter(torch.from_numpy(params[n].copy()).cuda()) for n, _, _ in cases.OPTIM_GROUPS}
    return tp, [{"params": [tp[n]], "lr": lr, "name": n} for n, _, lr in cases.OPTIM_GROUPS]


def test_adam_and_stats_match_goldens_and_oracle(golden_dir):
    gold = np.load(os.path.join(golden_dir, "optim.npz"))
    params, grads, radii, g2d = cases.optim_inputs()
    tp, l = _groups(params)
    opt = woptim.Adam(l, lr=0.0, eps=1e-15)     # scene/gaussian_model.py:181
    P = cases.OPTIM_P
    mr = torch.zeros((P,), device="cuda")
    acc, dn = torch.zeros((P, 1), device="cuda"), torch.zeros((P, 1), device="cuda")
    o = {n: (params[n], np.zeros_like(params[n]), np.zeros_like(params[n])) for n, _, _ in cases.OPTIM_GROUPS}
    omr, oacc, odn = (np.zeros(P, np.float32) for _ in range(3))
    for it in range(cases.OPTIM_STEPS):
        for g in opt.param_groups:
            if g["name"] == "xyz":
                g["lr"] = cases.OPTIM_XYZ_LRS[it]
        for n in tp:
            tp[n].grad = torch.from_numpy(grads[it][n]).cuda()
        woptim.add_densification_stats(torch.from_numpy(radii[it]).cuda(), torch.from_numpy(g2d[it]).cuda(), mr, acc, dn)
        opt.step()
        opt.zero_grad(set_to_none=True)
        omr, oacc, odn = optim_oracle.densify_stats(radii[it], g2d[it], omr, oacc, odn)
        assert np.array_equal(mr.cpu().numpy(), gold["step%d_max_radii2D" % it])
        assert np.array_equal(dn.cpu().numpy(), gold["step%d_denom" % it])
        assert np.array_equal(mr.cpu().numpy(), omr) and np.array_equal(dn.cpu().numpy().reshape(-1), odn)
        _close("xyz_gradient_accum", acc.cpu().numpy(), gold["step%d_xyz_gradient_accum" % it])
        _close("xyz_gradient_accum vs oracle", acc.cpu().numpy(), oacc)
        for n, _, lr in cases.OPTIM_GROUPS:
            lr = cases.OPTIM_XYZ_LRS[it] if n == "xyz" else lr
            o[n] = optim_oracle.adam_step(o[n][0], grads[it][n], o[n][1], o[n][2], it + 1, lr)
            st = opt.state[tp[n]]
            assert float(st["step"]) == it + 1
            for suffix, mine, orc in (("", tp[n].detach(), o[n][0]), ("_exp_avg", st["exp_avg"], o[n][1]),
                                      ("_exp_avg_sq", st["exp_avg_sq"], o[n][2])):
                _close("step %d %s%s" % (it, n, suffix), mine.cpu().numpy(), gold["step%d_%s%s" % (it, n, suffix)])
                _close("step %d %s%s vs oracle" % (it, n, suffix), mine.cpu().numpy(), orc)


def test_adam_is_bit_identical_to_torch_cuda_adam_at_training_size():
    """1 M Gaussians, the six groups of training_setup (59 M parameters), 3 steps: the fused kernel performs the
    float operations of torch's CUDA multi-tensor Adam in their order, so parameters and both moments agree
    bit for bit (if a torch build ever contracts differently this degrades to the optimiser tier)."""
    P = 1000000
    gen = torch.Generator(device="cuda").manual_seed(5)
    shapes = [(n, (P,) + s, lr) for n, s, lr in cases.OPTIM_GROUPS]
    init = {n: torch.randn(sh, device="cuda", generator=gen) for n, sh, _ in shapes}
    mine = {n: torch.nn.Parameter(init[n].clone()) for n in init}
    ref = {n: torch.nn.Parameter(init[n].clone()) for n in init}
    a = woptim.Adam([{"params": [mine[n]], "lr": lr, "name": n} for n, _, lr in shapes], lr=0.0, eps=1e-15)
    b = torch.optim.Adam([{"params": [ref[n]], "lr": lr, "name": n} for n, _, lr in shapes], lr=0.0, eps=1e-15)
    for it in range(3):
        for n, sh, _ in shapes:
            g = torch.randn(sh, device="cuda", generator=gen) * 1e-3
            g[::3] = 0.0   # invisible Gaussians
            mine[n].grad, ref[n].grad = g, g.clone()
        a.step()
        b.step()
        for n in init:
            for k in ("exp_avg", "exp_avg_sq"):
                x, y = a.state[mine[n]][k], b.state[ref[n]][k]
                if not torch.equal(x, y):
                    _close("%s %s step %d" % (n, k, it), x.cpu().numpy(), y.cpu().numpy())
            if not torch.equal(mine[n], ref[n]):
                _close("%s step %d" % (n, it), mine[n].detach().cpu().numpy(), ref[n].detach().cpu().numpy())
    exact = all(torch.equal(mine[n], ref[n]) for n in init)
    assert exact, "parameters are within tolerance of torch's CUDA Adam but not bit-identical"


def test_adam_keeps_the_reference_state_contract():
    """The reference edits optimizer.state[p]["exp_avg"/"exp_avg_sq"] and swaps group["params"][0] when it
    densifies or prunes (scene/gaussian_model.py:296-345); parameters without a gradient are skipped."""
    p = torch.nn.Parameter(torch.ones(10, 3, device="cuda"))
    q = torch.nn.Parameter(torch.ones(5, device="cuda"))
    opt = woptim.Adam([{"params": [p], "lr": 0.1, "name": "xyz"}, {"params": [q], "lr": 0.1, "name": "opacity"}],
                      lr=0.0, eps=1e-15)
    p.grad = torch.full_like(p, 0.5)
    opt.step()
    assert set(opt.state[p].keys()) == {"step", "exp_avg", "exp_avg_sq"} and q not in opt.state
    assert torch.allclose(p, torch.full_like(p, 0.9))   # first Adam step moves by lr * sign(g)
    mask = torch.arange(10, device="cuda") % 2 == 0
    group = opt.param_groups[0]
    stored = opt.state.get(group["params"][0], None)
    stored["exp_avg"] = stored["exp_avg"][mask]
    stored["exp_avg_sq"] = stored["exp_avg_sq"][mask]
    del opt.state[group["params"][0]]
    group["params"][0] = torch.nn.Parameter(group["params"][0][mask].requires_grad_(True))
    opt.state[group["params"][0]] = stored
    group["params"][0].grad = torch.full((5, 3), 0.5, device="cuda")
    opt.step()
    assert float(opt.state[group["params"][0]]["step"]) == 2
    sd = opt.state_dict()
    assert sd["param_groups"][0]["name"] == "xyz" and sd["param_groups"][0]["eps"] == 1e-15
    t = torch.optim.Adam([{"params": [torch.nn.Parameter(torch.ones(5, 3, device="cuda"))], "lr": 0.1, "name": "xyz"},
                          {"params": [torch.nn.Parameter(torch.ones(5, device="cuda"))], "lr": 0.1, "name": "opacity"}],
                         lr=0.0, eps=1e-15)
    t.load_state_dict(sd)   # checkpoints are interchangeable with torch.optim.Adam


def test_optim_rejects_cpu_tensors_and_handles_empty():
    p = torch.nn.Parameter(torch.ones(4))
    opt = woptim.Adam([p], lr=0.1)
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError):
        opt.step()
    e = torch.nn.Parameter(torch.ones(0, 3, device="cuda"))
    opt = woptim.Adam([e], lr=0.1)
    e.grad = torch.ones(0, 3, device="cuda")
    opt.step()
    z = torch.zeros(0, device="cuda")
    woptim.add_densification_stats(torch.zeros(0, dtype=torch.int32, device="cuda"), torch.zeros(0, 3, device="cuda"),
                                   z, z.clone(), z.clone())
    with pytest.raises(RuntimeError):
        woptim.add_densification_stats(torch.zeros(3, dtype=torch.int32), torch.zeros(3, 3), torch.zeros(3),
                                       torch.zeros(3, 1), torch.zeros(3, 1))
    with pytest.raises(ValueError):
        woptim.Adam([e], lr=0.1, amsgrad=True)


def test_adam_takes_gradient_slices_of_the_flat_bucket_at_odd_P():
    """P is arbitrary after densify_and_prune; the gradients the raw entry hands to the leaves are slices of ONE flat
    allocation at offsets 3P, 6P, ... (ops._backward), 16-byte aligned only when P % 4 == 0.  The fused step must
    take them and agree with torch.optim.Adam bit for bit."""
    from wheat_3dgs_b200 import optim as woptim
    for P in (20001, 20002, 20003):
        gen = torch.Generator(device="cuda").manual_seed(P)
        shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3),
                  "rotation": (P, 4)}
        init = {k: torch.randn(s, device="cuda", generator=gen) for k, s in shapes.items()}
        mine = {k: torch.nn.Parameter(v.clone()) for k, v in init.items()}
        ref = {k: torch.nn.Parameter(v.clone()) for k, v in init.items()}
        lrs = dict(xyz=0.00016, f_dc=0.0025, f_rest=0.000125, opacity=0.05, scaling=0.005, rotation=0.001)
        o1 = woptim.Adam([{"params": [mine[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
        o2 = torch.optim.Adam([{"params": [ref[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
        offs = {"xyz": 0, "f_dc": 6 * P, "f_rest": 9 * P, "opacity": 54 * P, "scaling": 55 * P, "rotation": 58 * P}
        misaligned = 0
        for _ in range(2):
            bucket = torch.randn(71 * P, device="cuda", generator=gen) * 1e-3
            for k, s in shapes.items():
                g = bucket[offs[k]:offs[k] + mine[k].numel()].view(s)
                misaligned += int(g.data_ptr() % 16 != 0)
                mine[k].grad = g
                ref[k].grad = g.clone()
            o1.step()
            o2.step()
        assert misaligned > 0
        for k in shapes:
            assert torch.equal(mine[k].detach(), ref[k].detach()), (P, k)
            assert int(o1.state[mine[k]]["step"].item()) == 2


def test_raw_training_loop_at_odd_P_and_bucket_guard():
    """forward_raw + optim.Adam at P % 4 != 0 (the advertised loop right after a densification), and the guard against a
    second backward into a bucket whose slices are still some parameter's .grad."""
    import wheat_3dgs_b200
    from wheat_3dgs_b200 import rasterizer as rz, scene, train
    P, W, H = 20001, 160, 120
    sc = scene.synthetic_scene(P, seed=0, scale_median=0.004)
    cam = scene.overhead_cameras(W, H)[17]
    dev = torch.device("cuda")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    cd = dict(cam, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]), campos=t(cam["campos"]))
    tr = train.Trainer(train.raw_params_from_scene(sc, dev), 3, dev)
    gt = torch.rand(3, H, W, device=dev)
    bg = torch.zeros(3, device=dev)
    before = tr.p["xyz"].detach().clone()
    l0 = float(tr.iteration(cd, gt, bg))
    l1 = float(tr.iteration(cd, gt, bg))
    assert np.isfinite(l0) and np.isfinite(l1) and not torch.equal(before, tr.p["xyz"].detach())
    # bucket guard
    bucket = torch.empty(71 * P, device=dev)
    rz.set_grad_bucket(bucket)
    try:
        m2d = torch.zeros_like(tr.p["xyz"], requires_grad=True)
        img = tr.render(cd, bg, m2d)[0]
        img.sum().backward()
        img2 = tr.render(cd, bg, m2d)[0]
        with pytest.raises(RuntimeError, match="alias the bucket"):
            img2.sum().backward()
    finally:
        rz.set_grad_bucket(None)


def test_trainer_loop_head_sets_the_position_rate_and_raises_the_sh_degree():
    """train_vanilla_3dgs.py:58-62 through Trainer.iteration(..., iteration=i): the xyz step of iteration i uses that
    iteration's scheduled rate (a step with the rate at 30000 moves xyz ~100x less than one with the rate at 1), and
    iteration 1000 renders with one more SH band."""
    from wheat_3dgs_b200 import scene, train
    P, W, H = 20000, 160, 120
    sc = scene.synthetic_scene(P, seed=0, scale_median=0.004)
    cam = scene.overhead_cameras(W, H)[17]
    dev = torch.device("cuda")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    cd = dict(cam, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]), campos=t(cam["campos"]))
    gt = torch.rand(3, H, W, device=dev)
    bg = torch.zeros(3, device=dev)
    moved = {}
    for it in (1, 30000):
        tr = train.Trainer(train.raw_params_from_scene(sc, dev), 2, dev)
        before = tr.p["xyz"].detach().clone()
        assert np.isfinite(float(tr.iteration(cd, gt, bg, iteration=it)))
        assert tr.get_lr("xyz") == train.expon_lr(it, 0.00016, 0.0000016, max_steps=30000)
        assert tr.active_sh_degree == (3 if it % 1000 == 0 else 2)
        moved[it] = float((tr.p["xyz"].detach() - before).abs().max())
    # first Adam step: |delta| = lr for every element with a non-zero gradient
    # (float32 positions of magnitude ~1 resolve a 1.6e-6 step to within a few per cent)
    assert abs(moved[1] / 0.0001599754409773643 - 1) < 1e-3 and abs(moved[30000] / 1.6e-06 - 1) < 0.08
