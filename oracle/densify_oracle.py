"""TEST INFRASTRUCTURE ONLY (oracle): numpy restatement of GaussianModel.densify_and_prune and reset_opacity,
step by step as the reference runs them (NOT the fused plan of csrc/ws_densify.cu, so it checks that plan):

    densify_and_prune      scene/gaussian_model.py:440-458
      densify_and_clone    :420-438        densification_postfix / cat_tensors_to_optimizer  :354-390
      densify_and_split    :392-418        prune_points / _prune_optimizer                   :310-352
    build_rotation         utils/general_utils.py:78-99
    reset_opacity          scene/gaussian_model.py:234-237 (+ replace_tensor_to_optimizer :295-308)

Pinned by tests/golden/densify.npz, written by tests/golden/make_golden_densify.py from the UNMODIFIED reference
class (scene/gaussian_model.py imported from /root/reference, run on the CPU).  float32 throughout; exp / log are
numpy's, so values that pass through them agree with torch to an ulp or two, not bitwise.

State = dict of numpy arrays: the six parameters ("xyz" [P,3], "f_dc" [P,1,3], "f_rest" [P,K,3], "opacity" [P,1],
"scaling" [P,3], "rotation" [P,4]), their Adam moments ("m_<name>", "v_<name>"), "which_object" int32 [P,1],
"xyz_gradient_accum" [P,1], "denom" [P,1], "max_radii2D" [P].
"""
import numpy as np

NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
f32 = np.float32


def sigmoid(x):
    return (f32(1) / (f32(1) + np.exp(-x.astype(f32)))).astype(f32)


def build_rotation(r):
    r = r.astype(f32)
    norm = np.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = np.zeros((q.shape[0], 3, 3), f32)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - w * z)
    R[:, 0, 2] = 2 * (x * z + w * y)
    R[:, 1, 0] = 2 * (x * y + w * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - w * x)
    R[:, 2, 0] = 2 * (x * z - w * y)
    R[:, 2, 1] = 2 * (y * z + w * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def _cat(st, ext, src_rows, new_src):
    """cat_tensors_to_optimizer + densification_postfix: parameters and _which_object grow, moments grow by zeros,
    the statistics are re-created as zeros."""
    out = dict(st)
    for n in NAMES:
        out[n] = np.concatenate([st[n], ext[n]], 0)
        for mv in ("m_", "v_"):
            if mv + n in st:
                out[mv + n] = np.concatenate([st[mv + n], np.zeros_like(ext[n])], 0)
    out["which_object"] = np.concatenate([st["which_object"], ext["which_object"]], 0)
    P = out["xyz"].shape[0]
    out["xyz_gradient_accum"] = np.zeros((P, 1), f32)
    out["denom"] = np.zeros((P, 1), f32)
    out["max_radii2D"] = np.zeros((P,), f32)
    return out, np.concatenate([src_rows, new_src])


def _prune(st, mask, src_rows):
    keep = ~mask
    out = {k: v[keep] for k, v in st.items()}
    return out, src_rows[keep]


def densify_and_prune(st, samples, max_grad, min_opacity, extent, max_screen_size, percent_dense):
    """Returns (new state, src_rows int32[rows out]: the source Gaussian of every output row, S).
    `samples` float32 [2S,3]: what the reference's torch.normal(mean=0, std=stds) returned (:399-400)."""
    st = {k: np.asarray(v) for k, v in st.items()}
    P0 = st["xyz"].shape[0]
    src = np.arange(P0, dtype=np.int32)
    with np.errstate(divide="ignore", invalid="ignore"):
        grads = (st["xyz_gradient_accum"].astype(f32) / st["denom"].astype(f32)).astype(f32)
    grads[np.isnan(grads)] = 0.0
    thr = f32(percent_dense * extent)
    # ---- clone (:420-438)
    scal = np.exp(st["scaling"].astype(f32))
    sel = (np.abs(grads[:, 0]) >= f32(max_grad)) & (scal.max(1) <= thr)
    ext = {n: st[n][sel] for n in NAMES}
    ext["which_object"] = st["which_object"][sel]
    st, src = _cat(st, ext, src, src[sel])
    # ---- split (:392-418)
    n_init = st["xyz"].shape[0]
    padded = np.zeros((n_init,), f32)
    padded[:P0] = grads[:, 0]
    scal = np.exp(st["scaling"].astype(f32))
    sel = (padded >= f32(max_grad)) & (scal.max(1) > thr)
    S = int(sel.sum())
    samples = np.asarray(samples, f32).reshape(2 * S, 3)
    rots = np.tile(build_rotation(st["rotation"][sel]), (2, 1, 1))
    new_xyz = (np.einsum("nij,nj->ni", rots, samples).astype(f32) + np.tile(st["xyz"][sel], (2, 1))).astype(f32)
    new_scaling = np.log((np.tile(scal[sel], (2, 1)) / f32(0.8 * 2)).astype(f32)).astype(f32)
    ext = {"xyz": new_xyz, "scaling": new_scaling, "rotation": np.tile(st["rotation"][sel], (2, 1)),
           "f_dc": np.tile(st["f_dc"][sel], (2, 1, 1)), "f_rest": np.tile(st["f_rest"][sel], (2, 1, 1)),
           "opacity": np.tile(st["opacity"][sel], (2, 1)), "which_object": np.tile(st["which_object"][sel], (2, 1))}
    st, src = _cat(st, ext, src, np.tile(src[sel], 2))
    st, src = _prune(st, np.concatenate([sel, np.zeros(2 * S, bool)]), src)
    # ---- prune (:444-450)
    mask = sigmoid(st["opacity"])[:, 0] < f32(min_opacity)
    if max_screen_size:
        big_vs = st["max_radii2D"] > f32(max_screen_size)
        big_ws = np.exp(st["scaling"].astype(f32)).max(1) > f32(0.1 * extent)
        mask = mask | big_vs | big_ws
    st, src = _prune(st, mask, src)
    return st, src, S


def reset_opacity(opacity):
    o = np.minimum(sigmoid(opacity), f32(0.01))
    return np.log(o / (f32(1) - o)).astype(f32)
