"""The on-disk Gaussian format the reference's scripts exchange: GaussianModel.save_ply /
load_ply (scene/gaussian_model.py:196-293), i.e. a binary little-endian PLY with one float32 property per
attribute: x y z nx ny nz f_dc_0..2 f_rest_0..(3(M-1)-1) opacity scale_0..2 rot_0..3 which_object.

The reference goes through the `plyfile` package (absent in this image), so this module reads and writes the
format directly with numpy.  PARITY UNPINNED: no reference-written file is available here to compare bytes with;
the tests check the header text `plyfile` emits for such an element, the attribute order and transposes of
save_ply (:212-228), round trips, and tolerance to reordered / missing optional properties as load_ply has (:233-279).
Host-side IO, not part of the hot path.
"""
import numpy as np


def attribute_names(n_dc, n_rest, n_scale=3, n_rot=4):
    """construct_list_of_attributes (scene/gaussian_model.py:196-210)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += ["f_dc_%d" % i for i in range(n_dc)]
    names += ["f_rest_%d" % i for i in range(n_rest)]
    names.append("opacity")
    names += ["scale_%d" % i for i in range(n_scale)]
    names += ["rot_%d" % i for i in range(n_rot)]
    names.append("which_object")
    return names


def _np(t):
    return t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)


def save_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation, which_object=None):
    """Raw parameters as GaussianModel holds them: xyz [P,3], features_dc [P,1,3], features_rest [P,M-1,3],
    opacity [P,1], scaling [P,3], rotation [P,4], which_object [P,1] (zeros if None)."""
    xyz = _np(xyz).astype(np.float32)
    P = xyz.shape[0]
    f_dc = np.ascontiguousarray(_np(features_dc).astype(np.float32).transpose(0, 2, 1)).reshape(P, -1)
    f_rest = np.ascontiguousarray(_np(features_rest).astype(np.float32).transpose(0, 2, 1)).reshape(P, -1)
    wo = np.zeros((P, 1), np.float32) if which_object is None else _np(which_object).astype(np.float32).reshape(P, 1)
    attrs = np.concatenate((xyz, np.zeros_like(xyz), f_dc, f_rest, _np(opacity).astype(np.float32).reshape(P, 1),
                            _np(scaling).astype(np.float32), _np(rotation).astype(np.float32), wo), axis=1)
    names = attribute_names(f_dc.shape[1], f_rest.shape[1], _np(scaling).shape[1], _np(rotation).shape[1])
    assert attrs.shape[1] == len(names)
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join("property float %s\n" % n for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(attrs, dtype="<f4").tobytes())


_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2",
              "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4",
              "double": "f8", "float64": "f8"}


def read_vertex_element(path):
    """The first element of a PLY file as a numpy structured array (binary little/big endian or ascii)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("%s is not a PLY file" % path)
        fmt, count, props, in_first, n_elements = None, 0, [], False, 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError("unexpected end of PLY header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                n_elements += 1
                in_first = n_elements == 1
                if in_first:
                    count = int(tok[2])
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError("list properties are not supported in the vertex element")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt == "ascii":
            data = np.loadtxt(f, max_rows=count, ndmin=2)
            out = np.empty(count, dtype=[(n, t) for n, t in props])
            for k, (n, _) in enumerate(props):
                out[n] = data[:, k]
            return out
        order = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, order + t) for n, t in props])
        return np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)


def load_ply(path, max_sh_degree=3):
    """load_ply (scene/gaussian_model.py:233-293) as numpy float32 arrays in GaussianModel's layouts: dict with xyz
    [P,3], features_dc [P,1,3], features_rest [P,M-1,3], opacity [P,1], scaling [P,3], rotation [P,4],
    which_object int32 [P,1] (zeros when the file has none)."""
    v = read_vertex_element(path)
    names = v.dtype.names
    P = v.shape[0]
    col = lambda n: np.asarray(v[n], dtype=np.float32)
    by_index = lambda prefix: sorted([n for n in names if n.startswith(prefix)], key=lambda n: int(n.split("_")[-1]))
    xyz = np.stack((col("x"), col("y"), col("z")), axis=1)
    f_dc = np.stack((col("f_dc_0"), col("f_dc_1"), col("f_dc_2")), axis=1).reshape(P, 3, 1)
    extra = by_index("f_rest_")
    if len(extra) != 3 * (max_sh_degree + 1) ** 2 - 3:   # the reference asserts the same (:249)
        raise ValueError("%d f_rest_ properties, expected %d for SH degree %d"
                         % (len(extra), 3 * (max_sh_degree + 1) ** 2 - 3, max_sh_degree))
    f_rest = np.stack([col(n) for n in extra], axis=1).reshape(P, 3, (max_sh_degree + 1) ** 2 - 1) if extra else \
        np.zeros((P, 3, 0), np.float32)
    scales = np.stack([col(n) for n in by_index("scale_")], axis=1)
    rots = np.stack([col(n) for n in by_index("rot")], axis=1)
    wo = np.asarray(v["which_object"]).astype(np.int32).reshape(P, 1) if "which_object" in names else \
        np.zeros((P, 1), np.int32)
    return {"xyz": xyz, "features_dc": np.ascontiguousarray(f_dc.transpose(0, 2, 1)),
            "features_rest": np.ascontiguousarray(f_rest.transpose(0, 2, 1)), "opacity": col("opacity").reshape(P, 1),
            "scaling": scales, "rotation": rots, "which_object": wo}
