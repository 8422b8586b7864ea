"""3DGS point-cloud PLY files without `plyfile`: the reader/writer pair of the reference's
gaussian_splatting/scene/gaussian_model.py:191-256 (save_ply / load_ply) as one structured-dtype numpy read or write of
the 62-float records (x y z, nx ny nz, f_dc_0..2, f_rest_0..44, opacity, scale_0..2, rot_0..3 for SH degree 3), plus the
xyz-only update main.py:315-320 does per dumped frame (there by deep-copying the whole model)."""
import os

import numpy as np

_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1", "char": "i1",
              "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2", "int": "<i4", "int32": "<i4",
              "uint": "<u4", "uint32": "<u4"}


def read_ply_vertices(path):
    """the `vertex` element of a binary_little_endian PLY as a structured numpy array (memory-mapped file read)"""
    with open(path, "rb") as f:
        head = b""
        while not head.endswith(b"end_header\n"):
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: no end_header")
            head += line
            if len(head) > 1 << 20:
                raise ValueError(f"{path}: header too long")
        offset = f.tell()
    lines = head.decode("ascii", "replace").split("\n")
    if lines[0].strip() != "ply":
        raise ValueError(f"{path}: not a PLY file")
    fmt = [l.split() for l in lines if l.startswith("format")]
    if not fmt or fmt[0][1] != "binary_little_endian":
        raise ValueError(f"{path}: only binary_little_endian PLY is supported (got {fmt})")
    count, props, in_vertex, before = None, [], False, 0
    for l in lines:
        t = l.split()
        if not t:
            continue
        if t[0] == "element":
            if in_vertex:
                break
            if t[1] == "vertex":
                in_vertex, count = True, int(t[2])
            elif count is None:
                raise ValueError(f"{path}: an element precedes `vertex`; not a 3DGS point cloud")
        elif t[0] == "property" and in_vertex:
            if t[1] == "list":
                raise ValueError(f"{path}: list property in the vertex element")
            props.append((t[2], _PLY_TYPES[t[1]]))
    if count is None:
        raise ValueError(f"{path}: no vertex element")
    dt = np.dtype(props)
    return np.fromfile(path, dtype=dt, count=count, offset=offset)


def _sorted_cols(v, prefix):
    names = sorted((n for n in v.dtype.names if n.startswith(prefix)), key=lambda x: int(x.split("_")[-1]))
    return names


def read_gaussian_ply(path, max_sh_degree=3):
    """dict of float32 arrays shaped as GaussianModel.load_ply builds its parameters (gaussian_model.py:216-256):
    xyz [P,3], features_dc [P,1,3], features_rest [P,(deg+1)^2-1,3], opacity [P,1], scaling [P,3], rotation [P,4]"""
    v = read_ply_vertices(path)
    P = v.shape[0]
    col = lambda names: np.stack([v[n].astype(np.float32) for n in names], axis=1) if names else np.zeros((P, 0), np.float32)
    xyz = col(["x", "y", "z"])
    f_dc = col(["f_dc_0", "f_dc_1", "f_dc_2"]).reshape(P, 3, 1)
    rest_names = _sorted_cols(v, "f_rest_")
    if len(rest_names) != 3 * (max_sh_degree + 1) ** 2 - 3:
        raise AssertionError(f"{path}: {len(rest_names)} f_rest_* properties, expected {3 * (max_sh_degree + 1) ** 2 - 3}")   # :231
    f_rest = col(rest_names).reshape(P, 3, (max_sh_degree + 1) ** 2 - 1)
    return {"xyz": xyz, "features_dc": np.ascontiguousarray(f_dc.transpose(0, 2, 1)),
            "features_rest": np.ascontiguousarray(f_rest.transpose(0, 2, 1)), "opacity": col(["opacity"]),
            "scaling": col(_sorted_cols(v, "scale_")), "rotation": col(_sorted_cols(v, "rot"))}


def attribute_names(n_dc=3, n_rest=45, n_scale=3, n_rot=4):
    """gaussian_model.py:176-189 construct_list_of_attributes"""
    return (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(n_dc)] + [f"f_rest_{i}" for i in range(n_rest)] +
            ["opacity"] + [f"scale_{i}" for i in range(n_scale)] + [f"rot_{i}" for i in range(n_rot)])


def write_gaussian_ply(path, g):
    """gaussian_model.py:191-209 save_ply from the dict read_gaussian_ply returns (normals are zeros, as there)"""
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    P = g["xyz"].shape[0]
    flat = lambda a: a.transpose(0, 2, 1).reshape(P, a.shape[1] * a.shape[2])   # channel-major (:196-197)
    f_dc = flat(np.asarray(g["features_dc"], np.float32))
    f_rest = flat(np.asarray(g["features_rest"], np.float32))
    cols = np.concatenate([np.asarray(g["xyz"], np.float32), np.zeros((P, 3), np.float32), f_dc, f_rest,
                           np.asarray(g["opacity"], np.float32).reshape(P, 1), np.asarray(g["scaling"], np.float32),
                           np.asarray(g["rotation"], np.float32)], axis=1)
    names = attribute_names(f_dc.shape[1], f_rest.shape[1], g["scaling"].shape[1], g["rotation"].shape[1])
    assert cols.shape[1] == len(names)
    head = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P + "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(head.encode("ascii"))
        f.write(np.ascontiguousarray(cols, dtype="<f4").tobytes())


def write_deformed_ply(src_path, dst_path, mask, new_xyz):
    """main.py:315-320 --save_pcd: the source cloud with the simulated Gaussians' positions replaced, written without
    copying the model: the records are read once, three columns are patched, the file is written back"""
    v = read_ply_vertices(src_path).copy()
    mask = np.asarray(mask, bool)
    new_xyz = np.asarray(new_xyz, np.float32)
    if mask.shape[0] != v.shape[0] or new_xyz.shape != (int(mask.sum()), 3):
        raise ValueError("mask / new_xyz do not match the cloud")
    for k, n in enumerate(("x", "y", "z")):
        col = v[n].copy()
        col[mask] = new_xyz[:, k]
        v[n] = col
    os.makedirs(os.path.dirname(os.path.abspath(dst_path)), exist_ok=True)
    with open(src_path, "rb") as f:
        head = b""
        while not head.endswith(b"end_header\n"):
            head += f.readline()
    with open(dst_path, "wb") as f:
        f.write(head)
        f.write(v.tobytes())
