"""COLMAP images.txt / cameras.txt -> stereo pose batch (/root/reference/main.py:169-217,
275-407, 602-632).  Closed-form checks plus tests/golden/colmap_ref.npz: the outputs of the reference's OWN
functions (cut out of main.py's syntax tree by tests/golden/make_colmap_golden.py and executed unmodified; PyGLM is
not installed, so its glm.* calls ran on a NumPy shim of GLM's float32 formulas)."""
import numpy as np
import pytest

from gaussiansplattingviewer_b200 import colmap, util

IMAGES = """# Image list with two lines of data per image:
#   IMAGE_ID, QW, QX, QY, QZ, TX, TY, TZ, CAMERA_ID, NAME
#   POINTS2D[] as (X, Y, POINT3D_ID)
# Number of images: 3
1 0.851773 0.0165051 0.503764 -0.142941 -0.737434 1.02973 3.74354 1 a.jpg
10.5 20.5 -1 30.1 40.2 7
2 1 0 0 0 0.5 -1.5 2.0 1 b.jpg

3 0.5 -0.5 0.5 0.5 3.0 0.25 -4.0 1 c.jpg
1 2 3
"""
CAMERAS = """# Camera list with one line of data per camera:
#   CAMERA_ID, MODEL, WIDTH, HEIGHT, PARAMS[]
1 PINHOLE 1160 522 1000.5 1001.5 580.0 261.0
"""


@pytest.fixture()
def model(tmp_path):
    (tmp_path / "images.txt").write_text(IMAGES)
    (tmp_path / "cameras.txt").write_text(CAMERAS)
    return tmp_path


def test_text_parsing_follows_the_reference_loop(model):
    rows = colmap.read_images_txt(str(model / "images.txt"))
    assert [r[0] for r in rows] == ["1", "2", "3"] and rows[2][-1] == "c.jpg" and all(len(r) == 10 for r in rows)
    cam = colmap.read_cameras_txt(str(model / "cameras.txt"))
    assert cam == {"id": 1, "model": "PINHOLE", "width": 1160, "height": 522, "fx": 1000.5, "fy": 1001.5, "cx": 580.0,
                   "cy": 261.0}
    (model / "bad.txt").write_text("1 1 0 0 0 0 0 0 1\n")
    with pytest.raises(ValueError):
        colmap.read_images_txt(str(model / "bad.txt"))


def test_quaternion_matrix_is_a_rotation():
    R = colmap.quaternion_to_rotation_matrix(0.3, -0.2, 0.9, 0.1)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-12) and np.isclose(np.linalg.det(R), 1.0)
    assert np.allclose(colmap.quaternion_to_rotation_matrix(2, 0, 0, 0), np.eye(3))
    with pytest.raises(ValueError):
        colmap.quaternion_to_rotation_matrix(0, 0, 0, 0)


def test_identity_pose_closed_form():
    # q = identity, T = (0.5,-1.5,2): camera at -T, flipped z axis -> looks along +z with up (0,-1,0)
    pl, pr = colmap.load_camera_positions("2 1 0 0 0 0.5 -1.5 2.0 1 b.jpg".split())
    eye = np.array([-0.5, 1.5, -2.0], np.float32)
    assert np.array_equal(pl["camera_position"], eye)
    V = pl["camera_view"]
    assert V.dtype == np.float32 and np.allclose(V @ np.append(eye, 1.0), [0, 0, 0, 1], atol=1e-6)
    assert np.allclose(V[:3, :3] @ np.array([0, 0, 1.0]), [0, 0, -1], atol=1e-6)   # forward +z maps to view -z
    assert np.allclose(V[:3, :3] @ np.array([0, -1.0, 0]), [0, 1, 0], atol=1e-6)   # world up (0,-1,0) maps to view +y
    assert np.array_equal(V, util.look_at(eye, eye + np.array([0, 0, 1.0]), [0, -1.0, 0]))


def test_right_pose_is_half_a_unit_to_the_right_with_equal_depths(model):
    vl, cl, vr, cr, ids = colmap.stereo_pose_batch(str(model))
    assert vl.shape == (3, 4, 4) and cr.shape == (3, 3) and list(ids) == [1, 2, 3]
    for i in range(3):
        assert np.allclose(vl[i][:3, :3] @ vl[i][:3, :3].T, np.eye(3), atol=1e-5)    # rigid
        assert np.array_equal(vl[i][1:], vr[i][1:]) and np.array_equal(vl[i][0, :3], vr[i][0, :3])
        assert vr[i][0, 3] == np.float32(vl[i][0, 3] + np.float32(-0.5))
        right_axis = vl[i][0, :3]                                                     # view +x in world space
        assert np.allclose(cr[i] - cl[i], 0.5 * right_axis, atol=1e-5)
        assert np.allclose(vr[i] @ np.append(cr[i], 1.0), [0, 0, 0, 1], atol=1e-5)
        p = np.array([0.3, -0.7, 1.9, 1.0], np.float32)
        assert (vl[i] @ p)[2] == (vr[i] @ p)[2]                                       # same view-space depth


def test_pose_math_matches_the_reference_functions_run_here(golden):
    """quaternion_to_rotation_matrix / create_look_at_from_colmap (main.py:167-217, pure NumPy): bit for bit.
    load_camera_positions (main.py:275-407): the reference's call sequence on the glm shim, view matrices to 1 ulp
    of their scale, the right camera position (glm.inverse there, -R^T t here) to 1e-6."""
    z = golden("colmap_ref.npz")
    rows = [str(r).split() for r in z["rows"]]
    assert [r[0] for r in rows] == ["1", "2", "3"]
    for i, e in enumerate(rows):
        qw, qx, qy, qz = (float(e[k]) for k in (1, 2, 3, 4))
        x, y, zz = (float(e[k]) for k in (5, 6, 7))
        assert np.array_equal(colmap.quaternion_to_rotation_matrix(qw, qx, qy, qz), z[f"rotmat_{i}"])
        pos, center, up = colmap.create_look_at_from_colmap(x, y, zz, qw, qx, qy, qz)
        assert np.array_equal(pos, z[f"pos_{i}"]) and np.array_equal(center, z[f"center_{i}"])
        assert np.array_equal(up, z[f"up_{i}"])
        pl, pr = colmap.load_camera_positions(e)
        assert np.array_equal(pl["camera_position"], z[f"campos_left_{i}"].astype(np.float32))
        assert np.array_equal(pl["camera_front"], z[f"front_{i}"])
        assert np.array_equal(pl["camera_up"], z[f"up_pose_{i}"].astype(np.float32))
        for got, want in ((pl["camera_view"], z[f"view_left_{i}"]), (pr["camera_view"], z[f"view_right_{i}"])):
            assert got.dtype == np.float32 and got.shape == (4, 4)
            ulp = np.spacing(np.maximum(np.abs(want), np.float32(1.0)).astype(np.float32))
            assert np.all(np.abs(got - want) <= ulp), (i, np.abs(got - want).max())
        assert np.allclose(pr["camera_position"], z[f"campos_right_{i}"], rtol=0, atol=1e-6)
        # the right view is the left one with x translation + baseline, bit for bit, also in the reference's result
        assert np.array_equal(z[f"view_right_{i}"][1:], z[f"view_left_{i}"][1:])
        assert z[f"view_right_{i}"][0, 3] == np.float32(z[f"view_left_{i}"][0, 3] + np.float32(-0.5))
