"""A hand-built Gaussian PLY fixture: `python tests/golden/make_golden_ply.py` writes tests/golden/point_cloud_fixture.ply
and tests/golden/point_cloud_fixture.json.

The file is assembled here byte by byte with `struct`, NOT with wheat_3dgs_b200.plyio and not with `plyfile` (absent
from this image): the header is the text a PLY 1.0 writer emits for one `vertex` element of float properties in the
order of GaussianModel.construct_list_of_attributes (scene/gaussian_model.py:196-210), and every row is packed as
little-endian float32 in the column order of save_ply's concatenation (:229: xyz, normals, f_dc, f_rest, opacity,
scale, rotation, which_object).  Values are small integers / dyadic fractions chosen so that each cell says where it
belongs:  features_rest[p][c][ch] = 1000 p + 10 c + ch  makes any transpose mistake visible.
The JSON holds the parameters in GaussianModel's own layouts ([P,1,3], [P,15,3] ...) as load_ply (:239-293) must
return them.  PARITY UNPINNED as far as plyfile's own bytes go (comment lines, property type spelling): this pins
the format and the transposes, not plyfile's writer.
"""
import json
import os
import struct

HERE = os.path.dirname(os.path.abspath(__file__))
P, DEG = 5, 3
M = (DEG + 1) ** 2


def params():
    out = {"xyz": [], "features_dc": [], "features_rest": [], "opacity": [], "scaling": [], "rotation": [],
           "which_object": []}
    for p in range(P):
        out["xyz"].append([p + 0.5, -(p + 0.25), 2.0 * p])
        out["features_dc"].append([[100.0 * p + ch for ch in range(3)]])                       # [1][3]
        out["features_rest"].append([[1000.0 * p + 10.0 * c + ch for ch in range(3)] for c in range(M - 1)])  # [15][3]
        out["opacity"].append([p - 2.0])
        out["scaling"].append([-(p + 1.0), -(p + 2.0), -(p + 3.0)])
        out["rotation"].append([1.0, 0.125 * p, -0.25 * p, 0.5 * p])
        out["which_object"].append([p % 3])
    return out


def main():
    pr = params()
    names = ["x", "y", "z", "nx", "ny", "nz"] + ["f_dc_%d" % i for i in range(3)] + \
        ["f_rest_%d" % i for i in range(3 * (M - 1))] + ["opacity"] + ["scale_%d" % i for i in range(3)] + \
        ["rot_%d" % i for i in range(4)] + ["which_object"]
    head = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    head += "".join("property float %s\n" % n for n in names) + "end_header\n"
    body = b""
    for p in range(P):
        row = list(pr["xyz"][p]) + [0.0, 0.0, 0.0]
        # save_ply: features.transpose(1, 2).flatten(start_dim=1) -> channel-major: f_rest_{ch * (M-1) + c}
        row += [pr["features_dc"][p][0][ch] for ch in range(3)]
        row += [pr["features_rest"][p][c][ch] for ch in range(3) for c in range(M - 1)]
        row += pr["opacity"][p] + pr["scaling"][p] + pr["rotation"][p] + [float(pr["which_object"][p][0])]
        assert len(row) == len(names)
        body += struct.pack("<%df" % len(row), *row)
    with open(os.path.join(HERE, "point_cloud_fixture.ply"), "wb") as f:
        f.write(head.encode("ascii") + body)
    json.dump({"P": P, "sh_degree": DEG, "params": pr}, open(os.path.join(HERE, "point_cloud_fixture.json"), "w"))
    print("wrote", len(head) + len(body), "bytes")


if __name__ == "__main__":
    main()
