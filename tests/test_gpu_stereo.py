"""Stereo capture batch (gsr_capture_stereo_async; /root/reference/main.py:844-917).

The capture is three renders + device packing with ONE shared depth order; the bar is that it is
BIT-IDENTICAL to the three independent frames the reference's loop draws (whose own parity with the oracle
is the job of tests/test_gpu_parity.py), that the uint16 packing equals NumPy's
``(disparity * 65535).astype(np.uint16)`` (main.py:878-879), and that the left image and the disparity
plane agree with the oracle within the image tolerance.
"""
import numpy as np
import pytest

from gaussiansplattingviewer_b200 import colmap, scenes, util
from oracle import oracle as O

pytestmark = pytest.mark.gpu

ROWS = ["1 0.98 0.05 -0.17 0.02 0.4 -0.3 9.0 1 a.jpg", "2 0.9 -0.1 0.4 0.1 -3.0 0.5 7.5 1 b.jpg",
        "3 1 0 0 0 0.0 0.0 10.0 1 c.jpg", "4 0.7 0.1 0.7 0.0 -8.0 0.3 1.0 1 d.jpg"]


def _renderer(g, W, H, mod=3):
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
    r = B200Renderer(W, H)
    r.update_gaussian_data(g)
    cam = util.Camera(H, W)
    r.update_camera_intrin(cam)
    r.set_render_mod(mod)
    return r, cam


def _independent(r, cam, pl, pr, mod):
    r.set_render_mod(mod)
    r.update_camera_pose(cam, True, pl)
    left = r.read_pixels_rgb8().copy()
    r.set_render_mod(-1)
    d = r.read_pixels_r32f().copy()
    disp = (d * 65535).astype(np.uint16)
    r.set_render_mod(mod)
    r.update_camera_pose(cam, True, pr)
    right = r.read_pixels_rgb8().copy()
    return left, disp, right, d


@pytest.mark.parametrize("share", [True, False])
@pytest.mark.parametrize("mod", [3, 0])
def test_capture_equals_three_independent_frames(share, mod):
    W, H = 464, 208
    g = scenes.synthetic_gaussians(60000, seed=4)
    r, cam = _renderer(g, W, H, mod)
    poses = [colmap.load_camera_positions(row.split()) for row in ROWS]
    outs = []
    for pl, pr in poses:                      # pipelined: two captures in flight
        o = r.capture_stereo(pl, pr, share_order=share)
        outs.append(o)
        if len(outs) >= 2:
            pass
    # buffers rotate over two slots: walk again, syncing per pose, and compare each with the independent frames
    r.sync_host()
    for pl, pr in poses:
        o = r.capture_stereo(pl, pr, share_order=share)
        r.sync_host()
        got = {k: v.copy() for k, v in o.items()}
        left, disp, right, _ = _independent(r, cam, pl, pr, mod)
        assert got["left"].shape == (H, W, 3) and got["disparity"].dtype == np.uint16
        assert np.array_equal(got["left"], left)
        assert np.array_equal(got["disparity"], disp)
        assert np.array_equal(got["right"], right)
        assert left.any() and right.any() and disp.any() and not np.array_equal(left, right)


def test_pipelined_slots_hold_the_right_pose():
    W, H = 320, 200
    g = scenes.synthetic_gaussians(30000, seed=8)
    r, cam = _renderer(g, W, H)
    poses = [colmap.load_camera_positions(row.split()) for row in ROWS]
    ref = []
    for pl, pr in poses:
        o = r.capture_stereo(pl, pr)
        r.sync_host()
        ref.append({k: v.copy() for k, v in o.items()})
    pending = []
    for i, (pl, pr) in enumerate(poses):
        o = r.capture_stereo(pl, pr)
        pending.append((i, o))
        if len(pending) == 2:                 # the capture submitted two calls ago is complete now? no: after the NEXT call
            pass
        if len(pending) == 3:
            j, oj = pending.pop(0)
            for k in oj:
                assert np.array_equal(oj[k], ref[j][k]), (j, k)
    r.sync_host()
    for j, oj in pending:
        for k in oj:
            assert np.array_equal(oj[k], ref[j][k]), (j, k)


def test_subset_and_uint16_wrap():
    W, H = 256, 144
    g = scenes.synthetic_gaussians(20000, seed=3)
    # a splat 0.2 in front of the camera (which sits at (0,0,-10) looking along +z) has a disparity > 1, so
    # disparity * 65535 exceeds 65535: the cast wraps like NumPy's
    g.xyz[0] = [0.0, 0.0, -9.8]
    g.scale[0] = [0.01, 0.01, 0.01]
    g.opacity[0] = 1.0
    r, cam = _renderer(g, W, H)
    pl, pr = colmap.load_camera_positions(ROWS[2].split())
    o = r.capture_stereo(pl, None, want=("disparity",))
    r.sync_host()
    assert set(o) == {"disparity"}
    got = o["disparity"].copy()
    r.set_render_mod(-1)
    r.update_camera_pose(cam, True, pl)
    d = r.read_pixels_r32f().copy()
    assert (d * 65535).max() > 65535.0
    assert np.array_equal(got, (d * 65535).astype(np.uint16))


def test_left_and_disparity_agree_with_the_oracle():
    W, H = 320, 200
    g = scenes.synthetic_gaussians(20000, seed=6)
    r, cam = _renderer(g, W, H)
    pl, pr = colmap.load_camera_positions(ROWS[0].split())
    o = r.capture_stereo(pl, pr)
    r.sync_host()
    for pose, key, mod in ((pl, "left", 3), (pr, "right", 3), (pl, "disparity", -1)):
        ocam = O.make_camera(pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal(),
                             pose["camera_position"], 1.0, mod, g.sh_dim)
        ref = O.render(g, ocam, W, H, clamp_unorm=(mod != -1))
        img = ref[0] if isinstance(ref, tuple) else ref
        if key == "disparity":
            want = img[..., 0] * 65535.0
            assert np.abs(o[key].astype(np.float64) - np.floor(want)).max() <= 65535.0 / 255.0
        else:
            want = np.clip(img[..., :3], 0, 1) * 255.0
            assert np.abs(o[key].astype(np.float64) - want).max() <= 1.0 + 0.5   # 1/255 + rounding


def test_overflow_is_reported_at_the_sync_point_and_repaired():
    W, H = 320, 200
    g = scenes.synthetic_gaussians(50000, seed=2)
    r, cam = _renderer(g, W, H)
    pl, pr = colmap.load_camera_positions(ROWS[2].split())
    o = r.capture_stereo(pl, pr)
    r.sync_host()
    ref = {k: v.copy() for k, v in o.items()}
    r.set_capacity(4096)
    with pytest.raises(RuntimeError):
        r.capture_stereo(pl, pr)
        r.sync_host()
    r.set_capacity(0)
    o = r.capture_stereo(pl, pr)              # buffers were grown at the sync point: submit again
    r.sync_host()
    for k in o:
        assert np.array_equal(o[k], ref[k]), k
