"""GaussianModel.densify_and_prune / reset_opacity on the device (csrc/ws_densify.cu through the C ABI) against
the golden output of the unmodified reference class and against the numpy oracle (oracle/densify_oracle.py)."""
import types

import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu
NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity",
        "scaling": "_scaling", "rotation": "_rotation"}
EXACT = ("f_dc", "f_rest", "opacity", "rotation")


def make_model(st, opt_cls=torch.optim.Adam, with_state=True):
    """An object with the reference GaussianModel's attributes (scene/gaussian_model.py:44-62, 167-181)."""
    dev = torch.device("cuda")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    m = types.SimpleNamespace()
    groups = []
    for n in NAMES:
        p = torch.nn.Parameter(t(st[n]).requires_grad_(True))
        setattr(m, ATTR[n], p)
        groups.append({"params": [p], "lr": 1e-3, "name": n})
    m.optimizer = opt_cls(groups, lr=0.0, eps=1e-15)
    if with_state:
        for g in m.optimizer.param_groups:
            n = g["name"]
            m.optimizer.state[g["params"][0]] = {"step": torch.tensor(1.0), "exp_avg": t(st["m_" + n]),
                                                 "exp_avg_sq": t(st["v_" + n])}
    m._which_object = t(st["which_object"])
    m.xyz_gradient_accum, m.denom, m.max_radii2D = t(st["xyz_gradient_accum"]), t(st["denom"]), t(st["max_radii2D"])
    m.percent_dense = cases.DENSIFY_TRAIN_ARGS["percent_dense"]
    return m


def state_of(m):
    st = {}
    for g in m.optimizer.param_groups:
        n, p = g["name"], g["params"][0]
        assert getattr(m, ATTR[n]) is p and p.requires_grad and isinstance(p, torch.nn.Parameter)
        st[n] = p.detach().cpu().numpy()
        s = m.optimizer.state.get(p)
        if s is not None:
            st["m_" + n], st["v_" + n] = s["exp_avg"].cpu().numpy(), s["exp_avg_sq"].cpu().numpy()
            assert float(s["step"]) == 1.0
    assert len(m.optimizer.state) in (0, 6)   # the old parameters' entries are gone
    st["which_object"] = m._which_object.cpu().numpy()
    for k in ("xyz_gradient_accum", "denom", "max_radii2D"):
        st[k] = getattr(m, k).cpu().numpy()
    return st


def compare(got, want, moments=True):
    rows = want["xyz"].shape[0]
    for n in NAMES:
        for pre in ("", "m_", "v_") if moments else ("",):
            a, b = got[pre + n], want[pre + n]
            assert a.shape == b.shape and a.dtype == b.dtype, (pre + n, a.shape, b.shape)
            if n in EXACT or pre:
                assert np.array_equal(a, b), pre + n
            else:   # split children: exp / log / 3x3 product, CUDA vs host libm (a few ulp; log-scales near 0 -> absolute)
                assert np.allclose(a, b, rtol=1e-6, atol=1e-6), (pre + n, np.abs(a - b).max())
    assert np.array_equal(got["which_object"], want["which_object"]) and got["which_object"].dtype == np.int32
    assert got["xyz_gradient_accum"].shape == (rows, 1) and got["denom"].shape == (rows, 1)
    assert got["max_radii2D"].shape == (rows,)
    assert not got["xyz_gradient_accum"].any() and not got["denom"].any() and not got["max_radii2D"].any()


def _gold(golden_dir):
    import os
    return np.load(os.path.join(golden_dir, "densify.npz"))


@pytest.mark.parametrize("tag", list(cases.DENSIFY_CASES))
def test_densify_matches_reference_class(golden_dir, tag):
    from wheat_3dgs_b200 import densify
    import densify_oracle
    gold = _gold(golden_dir)
    kw = cases.DENSIFY_CASES[tag]
    before = {k[len("before/"):]: gold[k] for k in gold.files if k.startswith("before/") and "/step_" not in k}
    m = make_model(before)
    samples = torch.from_numpy(gold[tag + "/samples"]).cuda()
    src, counts, _ = densify.densify_and_prune(m, kw["max_grad"], kw["min_opacity"], kw["extent"], kw["max_screen_size"],
                                               return_rows=True, samples=samples)
    want = {k[len(tag) + 1:]: gold[k] for k in gold.files if k.startswith(tag + "/")}
    compare(state_of(m), want)
    _, osrc, S = densify_oracle.densify_and_prune(before, gold[tag + "/samples"], kw["max_grad"], kw["min_opacity"],
                                                  kw["extent"], kw["max_screen_size"], m.percent_dense)
    assert np.array_equal(src.cpu().numpy(), osrc) and counts[3] == S
    assert counts[0] + counts[1] + 2 * counts[2] == want["xyz"].shape[0]


def _random_state(P, seed, K=15):
    inp = cases.densify_inputs(P=P, seed=seed, K=K)
    r = np.random.RandomState(seed + 1)
    st = {n: inp[n] for n in NAMES}
    for n in NAMES:
        st["m_" + n] = r.normal(0, 1e-3, inp[n].shape).astype(np.float32)
        st["v_" + n] = r.uniform(1e-8, 1e-5, inp[n].shape).astype(np.float32)
    for k in ("which_object", "xyz_gradient_accum", "denom", "max_radii2D"):
        st[k] = inp[k]
    return st


@pytest.mark.parametrize("screen", [20, None])
def test_densify_large_matches_oracle_with_own_draw(screen):
    """50 K Gaussians, SH degree 3 widths, samples drawn on the device by the same torch.normal call as the
    reference's: rows, order and values against the step-by-step numpy oracle; our fused Adam keeps working on the
    new parameters."""
    from wheat_3dgs_b200 import densify, optim
    import densify_oracle
    st = _random_state(50000, seed=21)
    m = make_model(st, opt_cls=optim.Adam)
    torch.manual_seed(3)
    src, counts, samples = densify.densify_and_prune(m, 0.0002, 0.005, 2.0, screen, return_rows=True)
    torch.manual_seed(3)   # the draw is the reference's: torch.normal(mean=zeros(2S,3), std=exp(scaling)[sel].repeat(2,1))
    sel = torch.from_numpy(((st["xyz_gradient_accum"] / np.where(st["denom"] == 0, np.nan, st["denom"]))[:, 0] >= np.float32(0.0002))
                           & (np.exp(st["scaling"]).max(1) > np.float32(0.02))).cuda()
    stds = torch.exp(torch.from_numpy(st["scaling"]).cuda())[sel].repeat(2, 1)
    again = torch.normal(mean=torch.zeros((stds.size(0), 3), device="cuda"), std=stds)
    assert torch.equal(again, samples)
    want, osrc, S = densify_oracle.densify_and_prune(st, samples.cpu().numpy(), 0.0002, 0.005, 2.0, screen, m.percent_dense)
    assert counts[3] == S and S > 1000 and counts[1] > 1000
    assert np.array_equal(src.cpu().numpy(), osrc)
    compare(state_of(m), want)
    for g in m.optimizer.param_groups:
        g["params"][0].grad = torch.full_like(g["params"][0], 1e-3)
    m.optimizer.step()
    assert float(m.optimizer.state[m._xyz]["step"]) == 2.0 and torch.isfinite(m._xyz).all()


def test_densify_nothing_selected_and_no_optimizer_state():
    """Gradients below the threshold: no clone, no split (no samples drawn), only the prune; parameters without
    optimizer state (before the first step) are handled as _prune_optimizer's else branch (:321-324)."""
    from wheat_3dgs_b200 import densify
    import densify_oracle
    st = _random_state(3000, seed=8, K=3)
    st["xyz_gradient_accum"] = np.zeros_like(st["xyz_gradient_accum"])
    m = make_model(st, with_state=False)
    src, counts, samples = densify.densify_and_prune(m, 0.0002, 0.005, 2.0, None, return_rows=True)
    assert counts[1] == 0 and counts[2] == 0 and counts[3] == 0 and samples is None
    want, osrc, S = densify_oracle.densify_and_prune(st, np.zeros((0, 3), np.float32), 0.0002, 0.005, 2.0, None, m.percent_dense)
    assert S == 0 and np.array_equal(src.cpu().numpy(), osrc) and 0 < len(osrc) < 3000
    compare(state_of(m), want, moments=False)


def test_reset_opacity_matches_reference_class(golden_dir):
    from wheat_3dgs_b200 import densify
    gold = _gold(golden_dir)
    before = {k[len("before/"):]: gold[k] for k in gold.files if k.startswith("before/") and "/step_" not in k}
    m = make_model(before)
    old = m._opacity
    densify.reset_opacity(m)
    assert m._opacity is not old and m.optimizer.param_groups[3]["params"][0] is m._opacity
    s = m.optimizer.state[m._opacity]
    assert old not in m.optimizer.state and not s["exp_avg"].any() and not s["exp_avg_sq"].any() and float(s["step"]) == 1.0
    assert np.allclose(m._opacity.detach().cpu().numpy(), gold["reset/opacity"], rtol=2e-6, atol=1e-6)
    ref = torch.log(torch.min(torch.sigmoid(old), torch.ones_like(old) * 0.01) / (1 - torch.min(torch.sigmoid(old), torch.ones_like(old) * 0.01)))
    assert torch.equal(m._opacity.detach(), ref.detach())   # bitwise against torch's own CUDA ops


def test_densify_errors():
    from wheat_3dgs_b200 import densify
    st = _random_state(100, seed=2, K=3)
    m = make_model(st)
    with pytest.raises(RuntimeError, match="max_grad must be > 0"):
        densify.densify_and_prune(m, 0.0, 0.005, 2.0, None)
    m.xyz_gradient_accum = m.xyz_gradient_accum.cpu()
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        densify.densify_and_prune(m, 0.0002, 0.005, 2.0, None)


def test_training_loop_survives_densification_and_opacity_reset():
    """train_vanilla_3dgs.py:65-112 in miniature on this library only: raw-parameter forward, fused loss, backward,
    statistics, fused Adam; after a few iterations densify_and_prune changes the number of Gaussians and reset_opacity
    swaps the opacity parameter; the loop carries on with the new parameters and their optimizer state."""
    import types as _t
    from wheat_3dgs_b200 import densify, loss as wloss, optim, scene
    from wheat_3dgs_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    P, W, H = 20000, 200, 150
    dev = torch.device("cuda")
    sc = scene.synthetic_scene(P, seed=3, n_heads=20, scale_median=0.01, extent=(0.3, 0.2, 0.12))
    cam = scene.overhead_cameras(W, H, extent=(0.3, 0.2, 0.12))[17]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
    op = np.clip(sc["opacities"], 1e-4, 1 - 1e-4)
    raw = {"xyz": sc["means3D"], "f_dc": sc["shs"][:, :1], "f_rest": sc["shs"][:, 1:], "opacity": np.log(op / (1 - op)),
           "scaling": np.log(sc["scales"]), "rotation": sc["rotations"]}
    m = _t.SimpleNamespace(percent_dense=0.01)
    groups = []
    for n, lr in (("xyz", 1.6e-4), ("f_dc", 2.5e-3), ("f_rest", 1.25e-4), ("opacity", 0.05), ("scaling", 5e-3), ("rotation", 1e-3)):
        p = torch.nn.Parameter(t(raw[n]).requires_grad_(True))
        setattr(m, ATTR[n], p)
        groups.append({"params": [p], "lr": lr, "name": n})
    m.optimizer = optim.Adam(groups, lr=0.0, eps=1e-15)
    m._which_object = torch.zeros((P, 1), dtype=torch.int32, device=dev)
    m.xyz_gradient_accum, m.denom = torch.zeros((P, 1), device=dev), torch.zeros((P, 1), device=dev)
    m.max_radii2D = torch.zeros((P,), device=dev)
    gt = torch.rand((3, H, W), device=dev)
    rs = GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=torch.zeros(3, device=dev),
        scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]), sh_degree=3,
        campos=t(cam["campos"]), prefiltered=False, debug=False)

    def iteration():
        means2D = torch.zeros_like(m._xyz, requires_grad=True)
        image, radii, depth, alpha = GaussianRasterizer(rs).forward_raw(m._xyz, means2D, m._features_dc, m._features_rest,
                                                                        m._opacity, m._scaling, m._rotation)
        loss = wloss.l1_ssim_loss(image, gt, 0.2)
        loss.backward()
        optim.add_densification_stats(radii, means2D.grad, m.max_radii2D, m.xyz_gradient_accum, m.denom)
        m.optimizer.step()
        m.optimizer.zero_grad(set_to_none=True)
        return float(loss.detach())

    first = [iteration() for _ in range(4)]
    assert all(np.isfinite(first)) and float(m.denom.max()) == 4.0
    densify.densify_and_prune(m, 1e-7, 0.005, 0.36, None)   # low threshold: every seen Gaussian is cloned or split
    P2 = m._xyz.shape[0]
    assert P2 > P and m._features_rest.shape == (P2, 15, 3) and m._which_object.shape == (P2, 1)
    assert m.optimizer.state[m._xyz]["exp_avg"].shape == (P2, 3) and float(m.optimizer.state[m._xyz]["step"]) == 4.0
    densify.reset_opacity(m)
    assert float(torch.sigmoid(m._opacity).max()) <= 0.0100001
    after = [iteration() for _ in range(3)]
    assert all(np.isfinite(after)) and float(m.optimizer.state[m._xyz]["step"]) == 7.0
    assert m.denom.shape == (P2, 1) and float(m.denom.max()) == 3.0
