"""Failure handling of the C ABI on the GPU: instance-buffer overflow on the asynchronous entry
points, more than 65536 tiles, a frozen order surviving gsr_sort_depth.  `pytest -m gpu`."""
import numpy as np
import pytest

from gaussiansplattingviewer_b200 import _gsr, scenes, util
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def R():
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
    return B200Renderer


def _renderer(R, g, W, H, mod=3):
    cam, pose = scenes.default_view(W, H)
    r = R(W, H)
    r.update_gaussian_data(g)
    r.set_render_mod(mod)
    r.update_camera_intrin(cam)
    r.update_camera_pose(cam, True, pose)
    return r, cam, pose


def test_draw_after_an_overflowed_frame_is_complete(R):
    """gsr_render cannot know that its frame overflows; the next draw() sees the sticky counter, the library
    grows the buffers and the frame is drawn again - never a silently black image for good."""
    import torch
    g = scenes.synthetic_gaussians(50000, seed=41)
    W, H = 640, 360
    r, _, _ = _renderer(R, g, W, H)
    want = r.draw().cpu().numpy().copy()
    assert want[..., 3].max() > 0.9
    r.set_capacity(2048)
    first = r.draw().cpu().numpy()          # overflows: incomplete by construction
    torch.cuda.synchronize()
    assert r.stats().overflow == 1 and not first.any()
    second = r.draw().cpu().numpy()         # repaired inside draw()
    assert np.array_equal(_bits(second), _bits(want))
    s = r.stats()
    assert s.overflow == 0 and s.capacity >= s.n_instances
    # the raw ABI call reports it instead of repairing silently
    r.set_capacity(2048)
    r.draw()
    torch.cuda.synchronize()
    import ctypes as C
    rc = r._lib.gsr_render(r._h, C.c_void_p(r._frame.data_ptr()), r._stream())
    assert rc == _gsr.GSR_EOVERFLOW
    assert np.array_equal(_bits(r.draw().cpu().numpy()), _bits(want))


def test_synchronous_overflow_does_not_poison_the_pipelined_path(R):
    g = scenes.synthetic_gaussians(40000, seed=42)
    W, H = 480, 270
    r, _, _ = _renderer(R, g, W, H)
    want = r.read_pixels_rgb8().copy()
    r.set_capacity(1024)
    got = r.read_pixels_rgb8().copy()       # overflows, grows, re-renders inside the call
    assert np.array_equal(got, want)
    bufs = [r.read_pixels_rgb8_async() for _ in range(3)]   # the third call is the first sync point
    r.sync_host()                                           # must not raise a stale GSR_EOVERFLOW
    for b in bufs:
        assert np.array_equal(b, want)
    cap = r.stats().capacity
    r.read_pixels_rgb8_async(); r.sync_host()
    assert r.stats().capacity == cap        # and nothing was grown needlessly


def test_more_than_65536_tiles(R):
    """17 tile bits (4096 x 4352 with 16-px tiles): k_tile_scan keeps out of the sort histograms, the tile sort
    counts its digits itself; image equal to the 32-px tiling (16 bits, histogram from the scan)."""
    g = scenes.synthetic_gaussians(60000, seed=43)
    W, H = 4096, 4352
    r, _, _ = _renderer(R, g, W, H)
    r.tune(_gsr.GSR_TUNE_TILE, 32)
    a = r.draw().cpu().numpy().copy()
    assert a[..., 3].max() > 0.5
    r.tune(_gsr.GSR_TUNE_TILE, 16)
    b = r.draw().cpu().numpy()
    s = r.stats()
    assert s.tiles_x * s.tiles_y == 256 * 272 > 65536 and not s.overflow
    # same splats in the same order per pixel; a finished pixel keeps blending until its warp is finished, and the
    # warps of a 16-px tile are not those of a 32-px tile: below 1e-4 per channel (tests/test_gpu_parity.py)
    assert np.abs(a - b).max() <= 2e-4
    b2 = r.draw().cpu().numpy()             # the forked (untimed) path twice: no stale histogram
    assert np.array_equal(_bits(b), _bits(b2))


def test_gaussian_order_keeps_a_frozen_order(R):
    g = scenes.synthetic_gaussians(30000, seed=44)
    W, H = 480, 270
    r, cam, pose = _renderer(R, g, W, H)
    r.set_auto_sort(False)
    r.sort_and_update(cam, True, pose)                     # freezes the order of `pose`
    _, pose2 = scenes.default_view(W, H, eye=(6.0, 1.0, -8.0))
    r.update_camera_pose(cam, True, pose2)
    frozen = r.draw().cpu().numpy().copy()
    order = r.gaussian_order(pose["camera_view"])          # borrows the sort buffers
    want, _ = O.sort_depth(g.xyz, pose["camera_view"])
    assert np.array_equal(order, want)
    again = r.draw().cpu().numpy()
    assert np.array_equal(_bits(again), _bits(frozen))     # still the frozen order, not a per-frame sort
    r.set_auto_sort(True)
    live = r.draw().cpu().numpy()
    assert not np.array_equal(_bits(live), _bits(frozen))


def test_graph_replay_is_bit_identical_to_direct_launches(R):
    """Frames are replayed from a captured CUDA graph from the third frame of a configuration on (GSR_TUNE_GRAPH);
    only the uniforms change between replays.  Same images, same launch counts as kernel-by-kernel enqueueing, across
    pose / scale / SH-degree changes (same graph), mode and resolution changes and a scene reload (new graphs)."""
    import torch
    g = scenes.synthetic_gaussians(60000, seed=45)
    g2 = scenes.synthetic_gaussians(45000, seed=46)
    views, pos = scenes.orbit_poses(6, radius=9.0)

    def run(graph_on):
        W, H = 640, 360
        r, cam, _ = _renderer(R, g, W, H)
        r.tune(_gsr.GSR_TUNE_GRAPH, graph_on)
        out = []

        def frames(n=6):
            for i in range(n):
                pose = {"camera_front": np.zeros(3, np.float32), "camera_up": np.zeros(3, np.float32),
                        "camera_position": pos[i % 6], "camera_view": views[i % 6]}
                r.update_camera_pose(cam, True, pose)
                out.append(r.draw().cpu().numpy().copy())
        frames()
        r.set_scale_modifier(1.7); r.set_render_mod(1); frames(4)          # uniforms only: same graph
        r.set_render_mod(-3); frames(4)                                      # other blend kernel
        r.set_render_mod(-1); frames(4)                                      # float target, 1.2x scale
        r.set_render_mod(3); r.set_scale_modifier(1.0)
        for _ in range(4):                                                   # pipelined read-back: two more graphs
            out.append(r.read_pixels_rgb8_async().copy() if False else None)
        out[-4:] = []
        bufs = [r.read_pixels_rgb8_async() for _ in range(5)]
        r.sync_host()
        out.extend(b.copy() for b in bufs[-2:])
        r.set_render_reso(480, 270)
        cam2 = util.Camera(270, 480)
        r.update_camera_intrin(cam2); frames(4)                              # new resolution: buffers reallocated
        r.update_gaussian_data(g2); frames(4)                                # new scene
        torch.cuda.synchronize()
        return out, r.launch_count()

    a, la = run(1)
    b, lb = run(0)
    assert len(a) == len(b) and la == lb
    for i, (x, y) in enumerate(zip(a, b)):
        assert x.shape == y.shape and np.array_equal(x.view(np.uint8), y.view(np.uint8)), i
    assert not np.array_equal(a[0], a[1])


def test_resident_scene_buffers_filled_in_place(R):
    """gsr_alloc_gaussians + gsr_scene_buffer: the receiving side of a scene broadcast
    (pose_shard.broadcast_scene_resident, bench.py at N > 1).  A handle whose resident planes are filled in place from
    another handle's planes (what dist.broadcast does across ranks) renders the same bits as one fed through
    update_gaussian_data; the planes alias the library's memory (no copy)."""
    g = scenes.synthetic_gaussians(50_000, seed=45)
    W, H = 480, 270
    src, cam, pose = _renderer(R, g, W, H)
    want = src.draw().cpu().numpy().copy()
    dst = R(W, H)
    dst.alloc_gaussians(len(g), g.sh_dim)
    a, b = src.scene_buffers(), dst.scene_buffers()
    assert [tuple(t.shape) for t in a] == [(len(g), 3), (len(g), 4), (len(g), 4), (len(g), g.sh_dim)]
    assert [tuple(t.shape) for t in b] == [tuple(t.shape) for t in a]
    assert np.array_equal(a[0].cpu().numpy(), g.xyz)                 # the planes ARE the scene
    assert np.array_equal(a[2].cpu().numpy()[:, 3:4], g.opacity)
    for ta, tb in zip(a, b):
        assert ta.data_ptr() != tb.data_ptr()
        tb.copy_(ta)                                                 # in place, like a broadcast
    dst.set_render_mod(3)
    dst.update_camera_intrin(cam)
    dst.update_camera_pose(cam, True, pose)
    got = dst.draw().cpu().numpy()
    assert np.array_equal(_bits(got), _bits(want))
    assert dst.stats().n_gaussians == len(g)
    again = dst.scene_buffers()
    assert [t.data_ptr() for t in again] == [t.data_ptr() for t in b]
