"""Host-side mirror of the reference's Python surface (util_gau, util.Camera) against values produced
by the reference's own functions (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest

from gaussiansplattingviewer_b200 import scenes, util, util_gau

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_naive_gaussian_matches_reference(golden):
    z = golden("naive_ref.npz")
    g, a, b = util_gau.naive_gaussian()
    assert (a, b) == (int(z["ret1"]), int(z["ret2"])) == (0, 0)
    for f in ["xyz", "rot", "scale", "opacity", "sh"]:
        got = getattr(g, f)
        assert got.dtype == z[f].dtype == np.float32 and got.shape == z[f].shape
        assert np.array_equal(got, z[f]), f
    assert np.array_equal(g.flat(), z["flat"]) and g.flat().flags["C_CONTIGUOUS"]
    assert len(g) == int(z["length"]) == 4 and g.sh_dim == int(z["sh_dim"]) == 3


def test_load_ply_matches_reference(golden):
    z = golden("ply_ref.npz")
    g, bbox, center = util_gau.load_ply(os.path.join(GOLDEN, "tiny.ply"))
    for f in ["xyz", "rot", "scale", "opacity", "sh"]:
        got = getattr(g, f)
        assert got.dtype == np.float32 and got.shape == z[f].shape, f
        assert np.array_equal(got.view(np.uint32), z[f].view(np.uint32)), f  # bit-exact activations
    assert g.sh.shape == (257, 48)
    assert np.array_equal(bbox, z["bbox"]) and np.array_equal(center, z["center"])
    assert np.allclose(np.linalg.norm(g.rot, axis=1), 1.0, atol=1e-6)
    assert (g.opacity > 0).all() and (g.opacity < 1).all() and (g.scale > 0).all()


def test_load_ply_rejects_wrong_sh_count(tmp_path):
    n = 5
    rng = np.random.default_rng(0)
    p = tmp_path / "bad.ply"
    util_gau.save_ply(str(p), rng.normal(size=(n, 3)), rng.normal(size=(n, 3)), rng.normal(size=(n, 24)),
                      rng.normal(size=n), rng.normal(size=(n, 3)), rng.normal(size=(n, 4)))
    with pytest.raises(AssertionError):  # /root/reference/util_gau.py:94
        util_gau.load_ply(str(p))


def test_ply_container_variants(tmp_path):
    # ascii and big-endian bodies parse to the same columns
    cols = np.arange(12, dtype=np.float32).reshape(4, 3)
    hdr = "ply\nformat {} 1.0\ncomment x\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n" \
          "element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    pa = tmp_path / "a.ply"
    pa.write_bytes(hdr.format("ascii").encode() + "\n".join(" ".join(str(v) for v in r) for r in cols).encode() + b"\n")
    pb = tmp_path / "b.ply"
    pb.write_bytes(hdr.format("binary_big_endian").encode() + cols.astype(">f4").tobytes())
    for p in (pa, pb):
        v = util_gau.read_ply_vertices(str(p))
        assert np.array_equal(np.stack([v["x"], v["y"], v["z"]], 1), cols)
    with pytest.raises(ValueError):
        (tmp_path / "c.ply").write_bytes(hdr.format("binary_little_endian").encode() + cols.tobytes()[:-4])
        util_gau.read_ply_vertices(str(tmp_path / "c.ply"))


def test_camera_intrinsics_match_reference(golden):
    z = golden("camera_ref.npz")
    for (h, w) in [(720, 1280), (1080, 1920), (2160, 3840), (522, 1160)]:
        c = util.Camera(h, w)
        P = c.get_project_matrix()
        assert P.dtype == np.float32 and np.array_equal(P, z[f"proj_{h}x{w}"])
        assert np.array_equal(np.array(c.get_htanfovxy_focal()), z[f"hff_{h}x{w}"])
        assert c.get_focal() == float(z[f"focal_{h}x{w}"])
    c = util.Camera(720, 1280)
    assert c.fovy == float(z["fovy"])
    assert np.array_equal(c.position, z["position"]) and np.array_equal(c.target, z["target"])
    assert np.array_equal(c.up, z["up"])
    # htany = 0.30314, htanx = 0.53892, focal = 1187.557 at 1280x720
    hx, hy, f = c.get_htanfovxy_focal()
    assert abs(hy - 0.30314) < 1e-5 and abs(hx - 0.53892) < 1e-5 and abs(f - 1187.557) < 1e-3


def test_look_at_is_right_handed_rigid():
    rng = np.random.default_rng(2)
    for _ in range(20):
        eye, tgt = rng.normal(size=3) * 5, rng.normal(size=3)
        V = util.look_at(eye, tgt, (0, -1, 0))
        R = V[:3, :3].astype(np.float64)
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-5) and np.linalg.det(R) > 0.999
        assert np.allclose(V @ np.append(eye, 1.0), [0, 0, 0, 1], atol=1e-4)  # eye -> origin
        t = V @ np.append(tgt, 1.0)
        assert t[2] < 0 and abs(t[0]) < 1e-4 and abs(t[1]) < 1e-4  # target straight ahead on -z
    # the three get_view_matrix branches of /root/reference/util.py:58-70
    c = util.Camera(720, 1280)
    assert np.allclose(c.get_view_matrix(False), util.look_at((0, 0, 3), (0, 0, 0), (0, -1, 0)))
    assert np.allclose(c.get_view_matrix(True), util.look_at((-3, 0, 1.5), (-3, 0, 0.5), (0, -1, 0)))
    pose = util.pose_from_lookat((1, 2, 3), (0, 0, -1), (0, -1, 0))
    assert np.array_equal(c.get_view_matrix(True, pose["camera_front"], pose["camera_position"], pose["camera_up"],
                                            pose["camera_view"]), pose["camera_view"])
    assert np.allclose(c.get_view_matrix(True, pose["camera_front"], pose["camera_position"], pose["camera_up"], None),
                       pose["camera_view"])


def test_synthetic_scene_is_seeded_and_well_formed():
    a = scenes.synthetic_gaussians(5000, seed=0)
    b = scenes.synthetic_gaussians(5000, seed=0)
    c = scenes.synthetic_gaussians(5000, seed=1)
    assert np.array_equal(a.flat(), b.flat()) and not np.array_equal(a.xyz, c.xyz)
    assert a.flat().shape == (5000, 59) and a.sh_dim == 48
    for f in ["xyz", "rot", "scale", "opacity", "sh"]:
        assert getattr(a, f).dtype == np.float32
    assert np.allclose(np.linalg.norm(a.rot, axis=1), 1, atol=1e-6)
    assert a.scale.min() >= 1e-3 * 0.999 and a.scale.max() <= 0.5 * 1.001


def test_trajectory_poses(golden):
    rows = np.load(os.path.join(GOLDEN, "camera_data.npy"))
    assert rows.shape == (18, 9)
    views, pos = scenes.trajectory_poses(rows, 64)
    assert views.shape == (64, 4, 4) and pos.shape == (64, 3) and views.dtype == np.float32
    for V in views:
        assert np.allclose(V[:3, :3] @ V[:3, :3].T, np.eye(3), atol=1e-4)
    # the loop passes through the (scaled) csv positions
    v18, p18 = scenes.trajectory_poses(rows, 18)
    assert np.allclose(p18, rows[:, 6:9] * 3.0, atol=1e-5)


def test_ply_row_layout_finds_columns_by_name():
    # the table the device decoder takes: offsets in PLY_COLUMNS order, whatever the file order is
    names = ["rot_3", "x", "nx", "z", "y"] + [f"f_rest_{i}" for i in reversed(range(45))] + \
            ["f_dc_2", "f_dc_0", "f_dc_1", "opacity", "scale_2", "scale_0", "scale_1", "rot_0", "rot_2", "rot_1"]
    props = [(n, "f4") for n in names]
    row_bytes, offs = util_gau.ply_row_layout(props)
    assert row_bytes == 4 * len(names) and len(offs) == 59 == len(util_gau.PLY_COLUMNS)
    for col, off in zip(util_gau.PLY_COLUMNS, offs):
        assert names[off // 4] == col
    with pytest.raises(AssertionError):  # /root/reference/util_gau.py:94
        util_gau.ply_row_layout([p for p in props if p[0] != "f_rest_7"])
    with pytest.raises(ValueError):
        util_gau.ply_row_layout([(n, "f8" if n == "opacity" else t) for n, t in props])
