"""Model persistence in the reference's two on-disk formats, so that a scene trained by the
reference loads into this engine and vice versa.

* ``.pt`` tensor set -- ``save_tensors`` / ``load_tensors`` (strategies/clm_offload/gaussian_model.py:236-290):
  ``xyz.pt opacity.pt scaling.pt rotation.pt parameters.pt`` (raw, pre-activation; parameters = [N,48] rows in the
  engine's [16 bases][3 rgb] order).
* INRIA-layout ``.ply`` -- ``save_ply`` / ``save_sub_plys`` / ``load_raw_ply`` / ``one_file_load_ply``
  (strategies/base_gaussian_model.py:165-247, clm_offload/gaussian_model.py:292-452): binary little-endian, one
  ``vertex`` element of 62 float32 properties ``x y z nx ny nz f_dc_0..2 f_rest_0..44 opacity scale_0..2 rot_0..3``.
  ``f_rest`` is CHANNEL-major on disk (f_rest_{c*15+k} = coefficient k+1 of colour channel c) while the engine's rows
  are basis-major (row[3 + 3k + c]); the transpose happens here.

The reference writes .ply through the ``plyfile`` package (not installed here); the files below are byte-compatible
with what it produces for this property list (same header lines, same packed little-endian float32 records).
"""
import os

import numpy as np
import torch


def ply_attribute_names(n_rest=45):
    """strategies/base_gaussian_model.py:165-187."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += ["f_dc_%d" % i for i in range(3)]
    names += ["f_rest_%d" % i for i in range(n_rest)]
    names += ["opacity"]
    names += ["scale_%d" % i for i in range(3)]
    names += ["rot_%d" % i for i in range(4)]
    return names


def rows_to_ply_features(rows):
    """[n,48] basis-major rows -> (f_dc [n,3], f_rest [n,45] channel-major) (base_gaussian_model.py:207-228)."""
    rows = np.asarray(rows, np.float32)
    n = rows.shape[0]
    f_dc = rows[:, :3]
    f_rest = rows[:, 3:].reshape(n, 15, 3).transpose(0, 2, 1).reshape(n, 45)
    return f_dc, f_rest


def ply_features_to_rows(f_dc, f_rest):
    """Inverse of rows_to_ply_features (gaussian_model.py:447-452: permute(0,2,1).reshape(N,-1))."""
    f_dc = np.asarray(f_dc, np.float32).reshape(-1, 3)
    n = f_dc.shape[0]
    f_rest = np.asarray(f_rest, np.float32).reshape(n, 3, -1)  # [n, channel, coeff]
    rest = f_rest.transpose(0, 2, 1).reshape(n, -1)
    return np.ascontiguousarray(np.concatenate([f_dc, rest], axis=1), np.float32)


def _write_ply(path, xyz, rows, opacity, scaling, rotation):
    n = xyz.shape[0]
    names = ply_attribute_names()
    rec = np.empty((n, len(names)), "<f4")
    f_dc, f_rest = rows_to_ply_features(rows)
    rec[:, 0:3] = xyz
    rec[:, 3:6] = 0.0  # normals
    rec[:, 6:9] = f_dc
    rec[:, 9:54] = f_rest
    rec[:, 54:55] = np.asarray(opacity, np.float32).reshape(n, 1)
    rec[:, 55:58] = scaling
    rec[:, 58:62] = rotation
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n
    header += "".join("property float %s\n" % a for a in names) + "end_header\n"
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(rec.tobytes())


_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
              "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def read_ply_vertices(path):
    """Minimal PLY reader for the vertex element: binary_little_endian or ascii, scalar properties in any order.
    -> numpy structured array."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("%s is not a PLY file" % path)
        fmt, n, props, in_vertex, seen_vertex = None, 0, [], False, False
        while True:
            line = f.readline()
            if not line:
                raise ValueError("%s: unterminated PLY header" % path)
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment" or tok[0] == "obj_info":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                if seen_vertex and tok[1] != "vertex":
                    in_vertex = False  # elements after the vertices are ignored
                else:
                    if tok[1] != "vertex":
                        raise ValueError("%s: first element is %r, expected vertex" % (path, tok[1]))
                    in_vertex, seen_vertex, n = True, True, int(tok[2])
            elif tok[0] == "property" and in_vertex:
                if tok[1] == "list":
                    raise ValueError("%s: list property in the vertex element" % path)
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        dt = np.dtype(props)
        if fmt == "binary_little_endian":
            data = np.frombuffer(f.read(n * dt.itemsize), dtype=dt, count=n)
        elif fmt == "ascii":
            flat = np.loadtxt(f, max_rows=n, ndmin=2)
            data = np.empty(n, dt)
            for i, (name, _) in enumerate(props):
                data[name] = flat[:, i]
        else:
            raise ValueError("%s: unsupported PLY format %r" % (path, fmt))
    if data.shape[0] != n:
        raise ValueError("%s: expected %d vertices, file holds %d" % (path, n, data.shape[0]))
    return data


def load_raw_ply(path, max_sh_degree=3):
    """gaussian_model.py:363-432 -> (xyz [N,3], features_dc [N,3,1], features_extra [N,3,15], opacities [N,1],
    scales [N,3], rots [N,4]) as float64 like the reference's np.zeros / np.stack path."""
    v = read_ply_vertices(path)
    names = v.dtype.names
    xyz = np.stack([np.asarray(v["x"]), np.asarray(v["y"]), np.asarray(v["z"])], axis=1).astype(np.float64)
    opacities = np.asarray(v["opacity"], np.float64)[..., None]
    features_dc = np.zeros((xyz.shape[0], 3, 1))
    for c in range(3):
        features_dc[:, c, 0] = v["f_dc_%d" % c]

    def sorted_group(prefix):
        return sorted([p for p in names if p.startswith(prefix)], key=lambda x: int(x.split("_")[-1]))

    extra = sorted_group("f_rest_")
    if len(extra) != 3 * (max_sh_degree + 1) ** 2 - 3:
        raise ValueError("%s holds %d f_rest properties, SH degree %d needs %d" %
                         (path, len(extra), max_sh_degree, 3 * (max_sh_degree + 1) ** 2 - 3))
    features_extra = np.stack([np.asarray(v[a], np.float64) for a in extra], axis=1) if extra else \
        np.zeros((xyz.shape[0], 0))
    features_extra = features_extra.reshape(xyz.shape[0], 3, (max_sh_degree + 1) ** 2 - 1)
    scales = np.stack([np.asarray(v[a], np.float64) for a in sorted_group("scale_")], axis=1)
    rots = np.stack([np.asarray(v[a], np.float64) for a in sorted_group("rot")], axis=1)
    return xyz, features_dc, features_extra, opacities, scales, rots


def _host(t):
    return t.detach().cpu().numpy()


def _current(gaussians):
    """The pinned SH table may lag behind (pipelined host Adam, lazy dense Adam): bring it up to date before reading."""
    if getattr(gaussians, "optimizer", None) is not None and hasattr(gaussians, "materialize_host_rows"):
        gaussians.materialize_host_rows()


def save_ply(gaussians, path):
    """strategies/base_gaussian_model.py:189-247."""
    _current(gaussians)
    _write_ply(path, _host(gaussians._xyz), gaussians._parameters.detach().numpy(), _host(gaussians._opacity),
               _host(gaussians._scaling), _host(gaussians._rotation))


def save_sub_plys(gaussians, path, n_split, split_size):
    """gaussian_model.py:292-361: ``<path>_rk<i>_ws<n_split>.ply`` holding rows [i*split_size, (i+1)*split_size)."""
    assert path.endswith(".ply")
    _current(gaussians)
    n = gaussians._xyz.shape[0]
    written = []
    for i in range(n_split):
        lo, hi = i * split_size, min((i + 1) * split_size, n)
        this = path[:-4] + "_rk" + str(i) + "_ws" + str(n_split) + ".ply"
        _write_ply(this, _host(gaussians._xyz[lo:hi]), gaussians._parameters[lo:hi].detach().numpy(),
                   _host(gaussians._opacity[lo:hi]), _host(gaussians._scaling[lo:hi]), _host(gaussians._rotation[lo:hi]))
        written.append(this)
    return written


def one_file_load_ply(gaussians, folder, capacity=None):
    """gaussian_model.py:434-487: ``<folder>/point_cloud.ply`` -> split GPU / pinned-host state."""
    xyz, f_dc, f_extra, opacities, scales, rots = load_raw_ply(os.path.join(folder, "point_cloud.ply"),
                                                               gaussians.max_sh_degree)
    rows = ply_features_to_rows(f_dc[:, :, 0], f_extra)
    gaussians.create_from_arrays(xyz.astype(np.float32), scales.astype(np.float32), rots.astype(np.float32),
                                 opacities.astype(np.float32), rows, capacity=capacity)
    gaussians.active_sh_degree = gaussians.max_sh_degree
    return gaussians


def save_tensors(gaussians, parent_path):
    """gaussian_model.py:236-244.  The SH rows are cloned first: saving the [:N] view would write the whole
    pre-allocated pinned buffer."""
    os.makedirs(parent_path, exist_ok=True)
    _current(gaussians)
    torch.save(gaussians._xyz.detach(), os.path.join(parent_path, "xyz.pt"))
    torch.save(gaussians._opacity.detach(), os.path.join(parent_path, "opacity.pt"))
    torch.save(gaussians._scaling.detach(), os.path.join(parent_path, "scaling.pt"))
    torch.save(gaussians._rotation.detach(), os.path.join(parent_path, "rotation.pt"))
    torch.save(gaussians._parameters.detach().clone(), os.path.join(parent_path, "parameters.pt"))


def load_tensors(gaussians, parent_path, capacity=None):
    """gaussian_model.py:246-290 (``capacity`` = the reference's ``args.prealloc_capacity``)."""
    def ld(name):
        t = torch.load(os.path.join(parent_path, name), map_location="cpu")
        return t.detach().float().contiguous().numpy()

    xyz, opacity, scaling, rotation, rows = ld("xyz.pt"), ld("opacity.pt"), ld("scaling.pt"), ld("rotation.pt"), \
        ld("parameters.pt")
    if capacity is not None and capacity < xyz.shape[0]:
        raise ValueError("capacity %d is smaller than the %d saved Gaussians" % (capacity, xyz.shape[0]))
    gaussians.create_from_arrays(xyz, scaling, rotation, opacity, rows, capacity=capacity)
    gaussians.active_sh_degree = gaussians.max_sh_degree
    return gaussians
