"""CPU: the Gaussian PLY format of GaussianModel.save_ply / load_ply (scene/gaussian_model.py:196-293).  Parity with
plyfile-written bytes is unpinned (plyfile is not installed here); these tests cover what the reference's code fixes:
attribute order, the [P,C,3] -> f_*_{c*?} transposes, round trips, optional / reordered properties."""
import json
import os

import numpy as np
import pytest

from wheat_3dgs_b200 import plyio


def _params(P=37, deg=3, seed=0):
    rng = np.random.default_rng(seed)
    M = (deg + 1) ** 2
    return dict(xyz=rng.normal(size=(P, 3)), features_dc=rng.normal(size=(P, 1, 3)),
                features_rest=rng.normal(size=(P, M - 1, 3)), opacity=rng.normal(size=(P, 1)),
                scaling=rng.normal(size=(P, 3)), rotation=rng.normal(size=(P, 4)),
                which_object=rng.integers(0, 9, size=(P, 1)))


@pytest.mark.parametrize("deg", [0, 1, 3])
def test_round_trip_and_layout(tmp_path, deg):
    p = _params(deg=deg)
    path = str(tmp_path / "point_cloud.ply")
    plyio.save_ply(path, **p)
    raw = open(path, "rb").read()
    head = raw[:raw.index(b"end_header\n") + 11].decode("ascii").splitlines()
    assert head[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 37"]
    names = [ln.split()[2] for ln in head if ln.startswith("property")]
    assert all(ln.split()[1] == "float" for ln in head if ln.startswith("property"))
    M = (deg + 1) ** 2
    assert names == plyio.attribute_names(3, 3 * (M - 1)) and names[:6] == ["x", "y", "z", "nx", "ny", "nz"]
    assert names[-1] == "which_object" and names[6:9] == ["f_dc_0", "f_dc_1", "f_dc_2"]
    assert len(raw) == raw.index(b"end_header\n") + 11 + 37 * 4 * len(names)
    v = plyio.read_vertex_element(path)
    # save_ply transposes [P, coeff, rgb] to channel-major before flattening (:218-219): f_rest_k = channel k // (M-1)
    if M > 1:
        assert np.array_equal(v["f_rest_1"], p["features_rest"].astype(np.float32)[:, 1, 0])
        assert np.array_equal(v["f_rest_%d" % (M - 1)], p["features_rest"].astype(np.float32)[:, 0, 1])
    assert np.all(v["nx"] == 0)
    back = plyio.load_ply(path, max_sh_degree=deg)
    for k in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"):
        assert back[k].dtype == np.float32 and np.array_equal(back[k], p[k].astype(np.float32)), k
    assert back["which_object"].dtype == np.int32 and np.array_equal(back["which_object"], p["which_object"])


def test_load_tolerates_reordered_missing_and_ascii(tmp_path):
    p = _params(P=5, deg=1)
    path = str(tmp_path / "a.ply")
    plyio.save_ply(path, **{k: v for k, v in p.items() if k != "which_object"})
    v = plyio.read_vertex_element(path)
    names = [n for n in v.dtype.names if n != "which_object"][::-1]          # reversed order, no which_object
    with open(str(tmp_path / "b.ply"), "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 5\n")
        f.write("".join("property float %s\n" % n for n in names) + "element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        for i in range(5):
            f.write(" ".join(repr(float(v[n][i])) for n in names) + "\n")
    back = plyio.load_ply(str(tmp_path / "b.ply"), max_sh_degree=1)
    assert np.array_equal(back["rotation"], p["rotation"].astype(np.float32))
    assert np.array_equal(back["features_rest"], p["features_rest"].astype(np.float32))
    assert np.array_equal(back["which_object"], np.zeros((5, 1), np.int32))
    with pytest.raises(ValueError):
        plyio.load_ply(path, max_sh_degree=3)                                 # the reference asserts the f_rest count


def test_hand_built_fixture_loads_and_resaves_byte_for_byte(tmp_path):
    """tests/golden/point_cloud_fixture.ply is packed with `struct` by tests/golden/make_golden_ply.py in the column
    order of save_ply (scene/gaussian_model.py:212-237): load_ply must return GaussianModel's layouts (:239-293, the
    f_rest reshape to [P,3,15] and transpose to [P,15,3]) and save_ply must reproduce the file byte for byte."""
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    path = os.path.join(gold, "point_cloud_fixture.ply")
    want = json.load(open(os.path.join(gold, "point_cloud_fixture.json")))
    got = plyio.load_ply(path, max_sh_degree=want["sh_degree"])
    for k, v in want["params"].items():
        ref = np.asarray(v, dtype=np.int32 if k == "which_object" else np.float32)
        assert got[k].shape == ref.shape and got[k].dtype == ref.dtype, k
        assert np.array_equal(got[k], ref), k
    out = str(tmp_path / "resaved.ply")
    plyio.save_ply(out, **got)
    assert open(out, "rb").read() == open(path, "rb").read()
