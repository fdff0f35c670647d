"""Persistence formats: INRIA-layout .ply and the .pt tensor set.  CPU only."""
import os

import numpy as np
import pytest
import torch

from clmgs import io as gio, model, scene as S


def _model(n=1500):
    spec = S.SceneSpec("t", n, 64, 64, 4, "object", 60.0, 0)
    g = S.make_gaussians(spec)
    gm = model.GaussianModelCLMOffload(3, "cpu").create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"],
                                                                   g["opacity_logit"], g["sh"], capacity=n + 100)
    return gm, g


def _same(a, b):
    for k in ("_xyz", "_opacity", "_scaling", "_rotation", "_parameters"):
        np.testing.assert_array_equal(getattr(a, k).numpy(), getattr(b, k).numpy(), err_msg=k)


def test_ply_layout_is_inria(tmp_path):
    gm, g = _model()
    p = str(tmp_path / "it" / "point_cloud.ply")
    gio.save_ply(gm, p)
    raw = open(p, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 1500"]
    props = [l.split()[-1] for l in lines[3:]]
    assert all(l.startswith("property float ") for l in lines[3:])
    assert props == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + \
        ["f_rest_%d" % i for i in range(45)] + ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
    rec = np.frombuffer(body, "<f4").reshape(1500, 62)
    np.testing.assert_array_equal(rec[:, 0:3], g["xyz"])
    assert not rec[:, 3:6].any()
    np.testing.assert_array_equal(rec[:, 54], g["opacity_logit"].reshape(-1))
    np.testing.assert_array_equal(rec[:, 55:58], g["log_scale"])
    np.testing.assert_array_equal(rec[:, 58:62], g["raw_quat"])
    # channel-major SH on disk: the reference's own expression (base_gaussian_model.py:207-228)
    rows = torch.from_numpy(np.asarray(g["sh"]))
    f_dc = rows[:, :3].contiguous().view(-1, 1, 3).transpose(1, 2).flatten(start_dim=1)
    f_rest = rows[:, 3:].contiguous().view(-1, 15, 3).transpose(1, 2).flatten(start_dim=1)
    np.testing.assert_array_equal(rec[:, 6:9], f_dc.numpy())
    np.testing.assert_array_equal(rec[:, 9:54], f_rest.numpy())
    # f_rest_{c*15+k} is coefficient k+1 of channel c
    assert rec[7, 9 + 1 * 15 + 4] == g["sh"][7, 3 + 3 * 4 + 1]


def test_ply_round_trip_and_raw_shapes(tmp_path):
    gm, g = _model()
    gio.save_ply(gm, str(tmp_path / "point_cloud.ply"))
    xyz, f_dc, f_extra, op, sc, rot = gio.load_raw_ply(str(tmp_path / "point_cloud.ply"))
    assert f_dc.shape == (1500, 3, 1) and f_extra.shape == (1500, 3, 15) and op.shape == (1500, 1)
    assert xyz.dtype == np.float64
    gm2 = gio.one_file_load_ply(model.GaussianModelCLMOffload(3, "cpu"), str(tmp_path), capacity=2000)
    _same(gm, gm2)
    assert gm2.parameters_buffer.shape == (2000, 48) and gm2.parameters_grad_buffer.shape == (2000, 48)


def test_sub_plys_cover_all_rows(tmp_path):
    gm, g = _model()
    files = gio.save_sub_plys(gm, str(tmp_path / "pc.ply"), 3, 600)
    assert [os.path.basename(f) for f in files] == ["pc_rk0_ws3.ply", "pc_rk1_ws3.ply", "pc_rk2_ws3.ply"]
    parts = [gio.read_ply_vertices(f) for f in files]
    assert [p.shape[0] for p in parts] == [600, 600, 300]
    np.testing.assert_array_equal(np.concatenate([p["x"] for p in parts]), g["xyz"][:, 0])
    np.testing.assert_array_equal(np.concatenate([p["f_rest_44"] for p in parts]), g["sh"][:, 3 + 3 * 14 + 2])


def test_reader_accepts_reordered_properties_ascii_and_doubles(tmp_path):
    names = gio.ply_attribute_names()
    rng = np.random.default_rng(0)
    vals = rng.standard_normal((5, len(names)))
    order = rng.permutation(len(names))
    p = tmp_path / "point_cloud.ply"
    with open(p, "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment made by a test\nelement vertex 5\n")
        for j in order:
            f.write("property %s %s\n" % ("double" if j % 2 else "float", names[j]))
        f.write("element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        for r in vals:
            f.write(" ".join(repr(float(r[j])) for j in order) + "\n")
    xyz, f_dc, f_extra, op, sc, rot = gio.load_raw_ply(str(p))
    np.testing.assert_allclose(xyz, vals[:, 0:3], rtol=1e-6)
    np.testing.assert_allclose(f_extra.reshape(5, 45), vals[:, 9:54], rtol=1e-6)
    np.testing.assert_allclose(rot, vals[:, 58:62], rtol=1e-6)


def test_reader_rejects_wrong_sh_degree_and_garbage(tmp_path):
    gm, g = _model(10)
    gio.save_ply(gm, str(tmp_path / "point_cloud.ply"))
    with pytest.raises(ValueError, match="f_rest"):
        gio.load_raw_ply(str(tmp_path / "point_cloud.ply"), max_sh_degree=2)
    (tmp_path / "x.ply").write_bytes(b"not a ply\n")
    with pytest.raises(ValueError):
        gio.read_ply_vertices(str(tmp_path / "x.ply"))
    raw = open(tmp_path / "point_cloud.ply", "rb").read()
    (tmp_path / "short.ply").write_bytes(raw[:-100])
    with pytest.raises(ValueError):
        gio.read_ply_vertices(str(tmp_path / "short.ply"))


def test_pt_tensor_set_round_trip(tmp_path):
    gm, g = _model()
    gio.save_tensors(gm, str(tmp_path / "ckpt"))
    assert sorted(os.listdir(tmp_path / "ckpt")) == ["opacity.pt", "parameters.pt", "rotation.pt", "scaling.pt", "xyz.pt"]
    # the view is cloned: the file holds N rows, not the pre-allocated capacity
    assert os.path.getsize(tmp_path / "ckpt" / "parameters.pt") < 1500 * 192 + 4096
    assert tuple(torch.load(tmp_path / "ckpt" / "parameters.pt").shape) == (1500, 48)
    gm2 = gio.load_tensors(model.GaussianModelCLMOffload(3, "cpu"), str(tmp_path / "ckpt"), capacity=1600)
    _same(gm, gm2)
    assert gm2.parameters_buffer.shape[0] == 1600
    with pytest.raises(ValueError):
        gio.load_tensors(model.GaussianModelCLMOffload(3, "cpu"), str(tmp_path / "ckpt"), capacity=10)
