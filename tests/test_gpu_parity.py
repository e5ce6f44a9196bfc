"""CUDA path (through the C ABI, via B200Renderer) against the oracle.  Needs a GPU: `pytest -m gpu`.

Bars: bit-exact for the per-Gaussian records, the sorted tile-instance list, the tile ranges and the
global depth order; images within 1/255 max-abs per channel and PSNR >= 50 dB (BASELINE.json
north_star).  Pixels whose value hangs on a threshold decision that is within 2e-5 relative of
flipping (the oracle flags them, see FRAG_EPS in oracle/gs_oracle.c) are exempt from the max-abs bar in
the hard-threshold modes and counted; the counts are reported (profiles/parity_report_*.txt).
"""
import numpy as np
import pytest

from gaussiansplattingviewer_b200 import _gsr, scenes, util, util_gau
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL = 1.0 / 255.0


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10 * np.log10(1.0 / mse)


def _report(line):
    """Observed parity figures go to stdout (pytest -s) and to gpurun_out/parity_report.txt (merged back from the
    GPU box; a copy of the last run is committed under profiles/)."""
    import os
    print(line)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        os.makedirs(os.path.join(root, "gpurun_out"), exist_ok=True)
        with open(os.path.join(root, "gpurun_out", "parity_report.txt"), "a") as f:
            f.write(line + "\n")
    except OSError:
        pass


def _check_image(img, ref, fragile, mod, tag):
    """The image bar of BASELINE.json: <= 1/255 max-abs per channel and PSNR >= 50 dB against the oracle.  In the
    hard-threshold modes (-3, -4) pixels the oracle flags as fragile (a 0.22 / 1/255 / power > 0 decision within 2e-5
    relative of flipping) are exempt; how many were exempt and how many pixels exceed the bar is reported."""
    diff = np.abs(img - ref).max(axis=-1)
    frag = fragile.astype(bool) if fragile is not None else np.zeros(diff.shape, bool)
    over = diff > TOL
    psnr = _psnr(img[..., :3], ref[..., :3])
    _report(f"[parity] {tag} mod {mod}: {diff.size} px, max|d| {diff.max():.3e}, over 1/255: {int(over.sum())} "
            f"(of them fragile {int((over & frag).sum())}), fragile px {int(frag.sum())}, PSNR {psnr:.1f} dB")
    assert frag.mean() < 2e-2  # sanity bound on the oracle's flags: 1.5 % at 6M / 3840x2160 (long lists), < 0.4 % elsewhere
    if mod in (-3, -4):
        assert not (over & ~frag).any(), (tag, mod, diff[~frag].max())
    else:
        assert not over.any(), (tag, mod, diff.max())
    assert psnr >= 50.0, (tag, mod, psnr)


@pytest.fixture(scope="module")
def R():
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
    return B200Renderer


def _setup(Rcls, g, W, H, eye=(0.0, 0.0, 10.0), mod=3, sm=1.0, keys64=False, debug=True):
    cam, pose = scenes.default_view(W, H, eye=eye)
    r = Rcls(W, H, keys64=keys64)
    r.debug = debug
    r.update_gaussian_data(g)
    r.set_scale_modifier(sm)
    r.set_render_mod(mod)
    r.update_camera_intrin(cam)
    r.update_camera_pose(cam, True, pose)
    ocam = O.make_camera(pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal(),
                         pose["camera_position"], sm, mod, g.sh_dim)
    return r, ocam


def _awkward_scene(n, seed):
    g = scenes.synthetic_gaussians(n, seed=seed)
    g.xyz[:8] = np.array([[0, 0, 12], [0, 0, 9.95], [7.0, 0, 0], [0, 4.0, 0], [0, 0, -95], [0, 0, 9.89], [0.1, 0.1, 5],
                          [-6.9, -3.9, 0]], dtype=np.float32)
    g.scale[8:12] = np.array([[3, 3, 3], [1e-4, 1e-4, 1e-4], [2.0, 1e-3, 1e-3], [0.5, 0.5, 1e-3]], dtype=np.float32)
    g.xyz[12] = g.xyz[13]  # exact depth tie
    return g


@pytest.mark.parametrize("mod,sm", [(3, 1.0), (2, 1.0), (1, 0.6), (0, 1.0), (-1, 1.0), (-2, 1.0), (-3, 2.0), (-4, 1.0)])
def test_preprocess_records_bit_exact(R, mod, sm):
    g = _awkward_scene(20000, seed=21)
    W, H = 640, 360
    r, ocam = _setup(R, g, W, H, mod=mod, sm=sm)
    r.draw()
    vs_gpu = r.debug_fetch(_gsr.GSR_DBG_VSOUT, np.float32, 16)
    st_gpu = r.debug_fetch(_gsr.GSR_DBG_SETUP, np.int32, 8)
    vs, st = O.preprocess(g, ocam, W, H)
    vs_ref = vs.view(np.float32).reshape(-1, 16)
    assert np.array_equal(vs_gpu[:, 0], vs_ref[:, 0])
    assert np.array_equal(_bits(vs_gpu[:, :14]), _bits(vs_ref[:, :14]))
    st_ref = st.view(np.int32).reshape(-1, 8)
    assert np.array_equal(st_gpu[:, 3], st_ref[:, 3])  # ok flag
    ok = st_ref[:, 3] == 1
    assert ok.sum() > 5000
    assert np.array_equal(st_gpu[ok], st_ref[ok])
    s = r.stats()
    assert s.n_visible == ok.sum() and s.n_gaussians == len(g) and not s.overflow


def test_sh_dim_3_scene_records_bit_exact(R):
    g, _, _ = util_gau.naive_gaussian()
    W, H = 1280, 720
    cam = util.Camera(H, W)
    r = R(W, H)
    r.debug = True
    r.update_gaussian_data(g)
    r.update_camera_intrin(cam)
    cam_pose = util.pose_from_lookat(cam.position, cam.target - cam.position, cam.up)
    r.update_camera_pose(cam, True, cam_pose)
    r.draw()
    vs_gpu = r.debug_fetch(_gsr.GSR_DBG_VSOUT, np.float32, 16)
    ocam = O.make_camera(cam_pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal(), cam.position,
                         1.0, 3, 3)
    vs, _ = O.preprocess(g, ocam, W, H)
    assert np.array_equal(_bits(vs_gpu[:, :14]), _bits(vs.view(np.float32).reshape(-1, 16)[:, :14]))


@pytest.mark.parametrize("tile", [32, 16])
@pytest.mark.parametrize("keys64", [False, True])
@pytest.mark.parametrize("W,H", [(640, 360), (333, 211)])
def test_instance_list_and_ranges_bit_exact(R, keys64, W, H, tile):
    g = _awkward_scene(30000, seed=22)
    r, ocam = _setup(R, g, W, H, keys64=keys64)
    r.tune(_gsr.GSR_TUNE_TILE, tile)
    r.draw()
    keys = r.debug_fetch(_gsr.GSR_DBG_INST_KEYS, np.uint64)
    vals = r.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32)
    ranges = r.debug_fetch(_gsr.GSR_DBG_RANGES, np.int32, 2)
    _, st = O.preprocess(g, ocam, W, H)
    okeys, ovals, oranges = O.tile_instances(st, W, H, tile)
    assert r.stats().n_instances == len(okeys) > 30000 * (16 / tile)
    assert np.array_equal(vals, ovals)  # stable order by (tile, depth, index)
    assert np.array_equal(keys, okeys)
    assert np.array_equal(ranges, oranges)


@pytest.mark.parametrize("mod,sm", [(3, 1.0), (0, 1.0), (-1, 1.0), (-2, 1.0), (-3, 1.0), (-4, 1.0), (3, 3.0)])
def test_images_within_tolerance(R, mod, sm):
    g = _awkward_scene(40000, seed=23)
    W, H = 800, 450
    r, ocam = _setup(R, g, W, H, eye=(1.0, -0.5, 9.0), mod=mod, sm=sm, debug=False)
    img = r.draw().cpu().numpy()
    ref, fragile, _ = O.render(g, ocam, W, H, clamp_unorm=(mod != -1), want_fragile=True)
    assert img.shape == ref.shape == (H, W, 4)
    assert ref[..., 3].max() > 0.9
    _check_image(img, ref, fragile, mod, f"40k/800x450 sm {sm}")


def test_config1_golden_images(R, golden):
    z = golden("naive_images_ref.npz")
    g, _, _ = util_gau.naive_gaussian()
    W, H = 1280, 720
    cam = util.Camera(H, W)
    pose = util.pose_from_lookat(cam.position, cam.target - cam.position, cam.up)
    assert np.array_equal(pose["camera_view"], z["view"])
    r = R(W, H)
    r.update_gaussian_data(g)
    r.update_camera_intrin(cam)
    r.update_camera_pose(cam, True, pose)
    for mod in [3, 0, -1, -2, -3, -4]:
        r.set_render_mod(mod)
        img = r.draw().cpu().numpy()
        want = z[f"img_mod{mod}"]
        d = np.abs(img - want)
        _report(f"[parity] config 1 golden (reference shaders) mod {mod}: max|d| {d.max():.3e}, over 1/255: "
                f"{int((d.max(axis=-1) > TOL).sum())} of {d.shape[0] * d.shape[1]} px")
        if mod in (-3, -4):
            assert (d.max(axis=-1) > TOL).mean() < 1e-5
        else:
            assert d.max() <= TOL, (mod, d.max())
        assert _psnr(img[..., :3], want[..., :3]) >= 50.0


def test_two_level_and_keys64_paths_agree_bitwise(R):
    g = scenes.synthetic_gaussians(200000, seed=24)
    W, H = 1920, 1080
    r1, _ = _setup(R, g, W, H, keys64=False, debug=False)
    r2, _ = _setup(R, g, W, H, keys64=True, debug=False)
    a = r1.draw().cpu().numpy()
    b = r2.draw().cpu().numpy()
    assert np.array_equal(_bits(a), _bits(b))
    v1 = r1.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32)
    v2 = r2.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32)
    assert np.array_equal(v1, v2) and len(v1) == r1.stats().n_instances


@pytest.mark.parametrize("kbits,n", [(32, 0), (32, 1), (32, 4095), (32, 4096), (32, 4097), (32, 1_000_003),
                                     (64, 1), (64, 3071), (64, 3073), (64, 700_001)])
def test_radix_sort_pairs_bit_exact(R, kbits, n):
    import torch
    r = R(64, 64)
    rng = np.random.default_rng(n + kbits)
    if kbits == 32:
        keys = rng.integers(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
        keys[: n // 3] &= np.uint32(0xFF)  # many ties
        tk = torch.from_numpy(keys.view(np.int32)).cuda()
    else:
        keys = rng.integers(0, 2 ** 63, size=n, dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=n, dtype=np.uint64)
        keys[: n // 3] &= np.uint64(0xFFFF)
        tk = torch.from_numpy(keys.view(np.int64)).cuda()
    vals = rng.permutation(n).astype(np.uint32)
    tv = torch.from_numpy(vals.view(np.int32)).cuda()
    r.sort_pairs(tk, tv)
    ok, ov = O.sort_pairs(keys, vals)
    assert np.array_equal(tk.cpu().numpy().view(keys.dtype), ok)
    assert np.array_equal(tv.cpu().numpy().view(np.uint32), ov)


def test_radix_sort_bit_ranges(R):
    import torch
    r = R(64, 64)
    rng = np.random.default_rng(5)
    n = 50_000
    keys = rng.integers(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    for b0, b1 in [(0, 13), (8, 24), (5, 6), (0, 32), (19, 32)]:
        tk = torch.from_numpy(keys.view(np.int32).copy()).cuda()
        tv = torch.from_numpy(vals.view(np.int32).copy()).cuda()
        r.sort_pairs(tk, tv, b0, b1)
        digit = (keys >> np.uint32(b0)) & np.uint32((1 << (b1 - b0)) - 1)
        want = np.argsort(digit, kind="stable")
        assert np.array_equal(tv.cpu().numpy().view(np.uint32), vals[want]), (b0, b1)
        assert np.array_equal(tk.cpu().numpy().view(np.uint32), keys[want]), (b0, b1)


def test_global_depth_order_matches_oracle_and_reference_golden(R, golden):
    z = golden("sort_ref.npz")
    cam = util.Camera(360, 640)
    for tag in ("a", "b"):
        xyz, view = z[f"xyz_{tag}"], z[f"view_{tag}"]
        n = len(xyz)
        g = util_gau.GaussianData(xyz, np.tile([1, 0, 0, 0], (n, 1)).astype(np.float32), np.ones((n, 3), np.float32),
                                  np.ones((n, 1), np.float32), np.zeros((n, 3), np.float32))
        r = R(640, 360)
        r.update_gaussian_data(g)
        got = r.gaussian_order(view)
        want, depth = O.sort_depth(xyz, view)
        assert got.dtype == np.int32 and got.shape == (n, 1)
        assert np.array_equal(got, want)
        # and the reference's own (unstable) result orders the same depths
        assert np.array_equal(depth[got[:, 0]], depth[z[f"index_{tag}"][:, 0]]) or tag == "a"
    # special values: NaN last, -0 == +0
    xyz = np.zeros((8, 3), np.float32)
    xyz[:, 2] = [np.nan, -0.0, 0.0, np.inf, -np.inf, 1e-40, -1e-40, -np.nan]
    g = util_gau.GaussianData(xyz, np.tile([1, 0, 0, 0], (8, 1)).astype(np.float32), np.ones((8, 3), np.float32),
                              np.ones((8, 1), np.float32), np.zeros((8, 3), np.float32))
    r = R(64, 64)
    r.update_gaussian_data(g)
    got = r.gaussian_order(np.eye(4, dtype=np.float32))
    assert np.array_equal(got[:, 0], np.argsort(xyz[:, 2], kind="stable"))
    del cam


def test_readback_variants_and_overflow_growth(R):
    g = scenes.synthetic_gaussians(50000, seed=25)
    W, H = 640, 360
    r, _ = _setup(R, g, W, H, debug=False)
    img = r.draw().cpu().numpy()
    rgb8 = r.read_pixels_rgb8().copy()
    assert rgb8.shape == (H, W, 3) and rgb8.dtype == np.uint8
    assert np.array_equal(rgb8, np.rint(np.clip(img[..., :3], 0, 1) * 255).astype(np.uint8))
    rgba = r.read_pixels_rgba32f().copy()
    assert np.array_equal(_bits(rgba), _bits(img))
    r.set_render_mod(-1)
    d = r.read_pixels_r32f().copy()
    img_d = r.draw().cpu().numpy()
    assert np.array_equal(_bits(d), _bits(img_d[..., 0])) and d.max() > 0
    # tiny capacity: the async path reports overflow, the host path grows and matches
    r.set_render_mod(3)
    r.set_capacity(2048)
    r.draw()
    s = r.stats()
    assert s.overflow == 1 and s.n_instances > 2048
    rgba2 = r.read_pixels_rgba32f().copy()
    assert np.array_equal(_bits(rgba2), _bits(img))
    assert r.stats().overflow == 0 and r.stats().capacity >= s.n_instances


def test_empty_and_degenerate_scenes(R):
    W, H = 320, 200
    cam, pose = scenes.default_view(W, H)
    # everything behind the camera -> black frame, zero instances
    g = scenes.synthetic_gaussians(1000, seed=1)
    g.xyz[:, 2] = 50.0
    r = R(W, H)
    r.update_gaussian_data(g)
    r.update_camera_intrin(cam)
    r.update_camera_pose(cam, True, pose)
    img = r.draw().cpu().numpy()
    assert not img.any() and r.stats().n_instances == 0 and r.stats().n_visible == 0
    # one splat covering the whole screen touches every tile
    g1 = util_gau.GaussianData(np.zeros((1, 3), np.float32), np.array([[1, 0, 0, 0]], np.float32),
                               np.full((1, 3), 5.0, np.float32), np.full((1, 1), 0.5, np.float32),
                               np.zeros((1, 48), np.float32))
    r.update_gaussian_data(g1)
    img = r.draw().cpu().numpy()
    s = r.stats()
    assert s.n_instances == s.tiles_x * s.tiles_y == 10 * 7   # 32x32 tiles by default
    assert img[..., 3].min() > 0.2 and abs(img[H // 2, W // 2, 3] - 0.5) < 0.01
    # NaN / inf inputs never crash and draw nothing
    g2 = util_gau.GaussianData(np.full((4, 3), np.nan, np.float32), np.zeros((4, 4), np.float32),
                               np.full((4, 3), np.inf, np.float32), np.ones((4, 1), np.float32),
                               np.zeros((4, 48), np.float32))
    r.update_gaussian_data(g2)
    assert not r.draw().cpu().numpy().any()


def test_call_order_errors(R):
    r = R(64, 64)
    with pytest.raises(RuntimeError):
        r.draw()
    cam, pose = scenes.default_view(64, 64)
    r.update_camera_intrin(cam)
    r.update_camera_pose(cam, True, pose)
    with pytest.raises(RuntimeError):  # no gaussians yet
        r.draw()
    with pytest.raises(RuntimeError):
        r.set_render_reso(0, 10)


def test_pose_batch_matches_single_draws(R):
    import torch
    g = scenes.synthetic_gaussians(30000, seed=26)
    W, H = 480, 270
    r, _ = _setup(R, g, W, H, debug=False)
    views, pos = scenes.orbit_poses(5)
    out = torch.empty((5, H, W, 4), dtype=torch.float32, device="cuda")
    r.draw_batch(views, pos, out)
    batch = out.cpu().numpy()
    cam = util.Camera(H, W)
    for i in range(5):
        pose = {"camera_front": np.zeros(3, np.float32), "camera_up": np.zeros(3, np.float32),
                "camera_position": pos[i], "camera_view": views[i]}
        r.update_camera_pose(cam, True, pose)
        single = r.draw().cpu().numpy()
        assert np.array_equal(_bits(single), _bits(batch[i])), i
    # and a frame of the batch against the oracle (not only the CUDA path against itself)
    i = 3
    ocam = O.make_camera(views[i], cam.get_project_matrix(), cam.get_htanfovxy_focal(), pos[i], 1.0, 3, g.sh_dim)
    ref, fragile, _ = O.render(g, ocam, W, H, clamp_unorm=True, want_fragile=True)
    _check_image(batch[i], ref, fragile, 3, "pose batch, 30k/480x270")


def test_pipelined_readback_matches_synchronous(R):
    """gsr_render_rgb8_host_async: frames land in their pinned buffers in submission order, bit-identical to
    the synchronous read-back, incl. after an overflow report."""
    g = scenes.synthetic_gaussians(40000, seed=27)
    W, H = 480, 270
    r, _ = _setup(R, g, W, H, debug=False)
    views, pos = scenes.orbit_poses(7)
    cam = util.Camera(H, W)

    def pose(i):
        return {"camera_front": np.zeros(3, np.float32), "camera_up": np.zeros(3, np.float32),
                "camera_position": pos[i], "camera_view": views[i]}

    want = []
    for i in range(7):
        r.update_camera_pose(cam, True, pose(i))
        want.append(r.read_pixels_rgb8().copy())
    got = []
    pending = []
    for i in range(7):
        r.update_camera_pose(cam, True, pose(i))
        pending.append(r.read_pixels_rgb8_async())
        if len(pending) == 3:  # the buffer of two calls ago is complete now
            got.append(pending.pop(0).copy())
    r.sync_host()
    got += [b.copy() for b in pending]
    assert len(got) == 7
    for i in range(7):
        assert np.array_equal(got[i], want[i]), i
    assert not np.array_equal(want[0], want[3])
    # overflow inside the pipeline is reported at a sync point, the buffers grow, a re-submit succeeds
    r.set_capacity(1024)
    r.update_camera_pose(cam, True, pose(2))
    r.read_pixels_rgb8_async()
    with pytest.raises(RuntimeError):
        r.sync_host()
    buf = r.read_pixels_rgb8_async()
    r.sync_host()
    assert np.array_equal(buf, want[2])


def test_full_size_properties_3m_1080p(R):
    """BASELINE config 3 size: properties that need no oracle run."""
    g = scenes.synthetic_gaussians(3_000_000, seed=0)
    W, H = 1920, 1080
    r, _ = _setup(R, g, W, H, debug=False)
    img = r.draw().cpu().numpy()
    s = r.stats()
    assert not s.overflow and s.n_visible > 1_000_000 and s.n_instances > s.n_visible
    keys = r.debug_fetch(_gsr.GSR_DBG_INST_KEYS, np.uint64)
    vals = r.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32)
    ranges = r.debug_fetch(_gsr.GSR_DBG_RANGES, np.int32, 2)
    assert len(keys) == s.n_instances
    assert np.all(keys[1:] >= keys[:-1])  # sortedness
    same = keys[1:] == keys[:-1]
    assert np.all(vals[1:][same] > vals[:-1][same])  # stability: ties keep ascending index
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    assert tiles.max() < s.tiles_x * s.tiles_y
    counts = np.bincount(tiles, minlength=s.tiles_x * s.tiles_y)
    assert np.array_equal(ranges[:, 1] - ranges[:, 0], counts)
    nz = counts > 0
    assert np.array_equal(ranges[nz, 0], (np.cumsum(counts) - counts)[nz])
    assert np.isfinite(img).all() and img[..., 3].max() <= 1.0 and img[..., :3].min() >= 0.0
    # idempotence: same state, same image, bit for bit
    assert np.array_equal(_bits(r.draw().cpu().numpy()), _bits(img))
    # and the image itself against the oracle at the real size (the C oracle needs a second or two on the host cores)
    cam, pose = scenes.default_view(W, H)
    ocam = O.make_camera(pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal(),
                         pose["camera_position"], 1.0, 3, g.sh_dim)
    ref, fragile, _ = O.render(g, ocam, W, H, clamp_unorm=True, want_fragile=True)
    _check_image(img, ref, fragile, 3, "config 3: 3M SH3 1920x1080")


def test_full_size_image_config2_1m_1080p(R):
    """BASELINE config 2: 1M SH3 at 1920x1080, one view, image against the oracle."""
    g = scenes.synthetic_gaussians(1_000_000, seed=2)
    W, H = 1920, 1080
    r, ocam = _setup(R, g, W, H, debug=False)
    img = r.draw().cpu().numpy()
    s = r.stats()
    assert not s.overflow and s.n_visible > 300_000
    ref, fragile, _ = O.render(g, ocam, W, H, clamp_unorm=True, want_fragile=True)
    _check_image(img, ref, fragile, 3, "config 2: 1M SH3 1920x1080")
    # per-Gaussian records at this size too: bit-exact
    r.debug = True
    r.draw()
    vs_gpu = r.debug_fetch(_gsr.GSR_DBG_VSOUT, np.float32, 16)
    vs, st = O.preprocess(g, ocam, W, H)
    assert np.array_equal(_bits(vs_gpu[:, :14]), _bits(vs.view(np.float32).reshape(-1, 16)[:, :14]))
    vals = r.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32)
    ranges = r.debug_fetch(_gsr.GSR_DBG_RANGES, np.int32, 2)
    _, ovals, oranges = O.tile_instances(st, W, H, 32)
    assert np.array_equal(vals, ovals) and np.array_equal(ranges, oranges)


@pytest.mark.parametrize("mod", [3, -1, -3])
def test_full_size_properties_6m_4k(R, mod):
    """BASELINE config 4 size (6M SH3, 3840x2160, modes 3 / -1 / -3): size-independent properties, and the
    one-sort 64-bit-key path as an independent cross-check of the two-level sort at full size."""
    g = scenes.synthetic_gaussians(6_000_000, seed=1)
    W, H = 3840, 2160
    r, _ = _setup(R, g, W, H, mod=mod, debug=False)
    img = r.read_pixels_rgba32f().copy()         # host path: grows the instance buffers if mode -1 needs it
    s = r.stats()
    assert not s.overflow and s.n_visible > 2_000_000 and s.n_instances > 2 * s.n_visible
    keys = r.debug_fetch(_gsr.GSR_DBG_INST_KEYS, np.uint64)
    vals = r.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32)
    ranges = r.debug_fetch(_gsr.GSR_DBG_RANGES, np.int32, 2)
    assert len(keys) == s.n_instances
    assert np.all(keys[1:] >= keys[:-1])
    same = keys[1:] == keys[:-1]
    assert np.all(vals[1:][same] > vals[:-1][same])
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    counts = np.bincount(tiles, minlength=s.tiles_x * s.tiles_y)
    assert s.tiles_x * s.tiles_y == 120 * 68 and np.array_equal(ranges[:, 1] - ranges[:, 0], counts)
    assert np.isfinite(img).all() and img[..., 3].max() <= 1.0
    if mod != -1:
        assert img[..., :3].min() >= 0.0 and img[..., :3].max() <= 1.0
    del keys, vals
    r2, ocam = _setup(R, g, W, H, mod=mod, keys64=True, debug=False)
    img2 = r2.read_pixels_rgba32f()
    assert np.array_equal(_bits(img2), _bits(img))
    del r2
    # the image against the oracle at the real size: colour, the 1.2x-scale disparity pass (gau_vert.glsl:151-156,
    # 182-211; float target, unclamped) and the flat-ball threshold mode
    ref, fragile, _ = O.render(g, ocam, W, H, clamp_unorm=(mod != -1), want_fragile=True)
    _check_image(img, ref, fragile, mod, "config 4: 6M SH3 3840x2160")


@pytest.mark.parametrize("mod,sm", [(3, 1.0), (0, 2.0), (-1, 1.0), (-2, 1.0), (-3, 1.0), (-4, 1.0)])
def test_tile_edge_16_and_32_give_identical_images(R, mod, sm):
    """The tile edge only decides how the instance list is cut up; every pixel sees the same splats in the same order.
    The hard-threshold modes are bit-identical.  In the gaussian modes a pixel keeps blending until its whole WARP is
    finished (T < 1e-4 everywhere; the shader path itself never stops), and which pixels share a warp depends on the
    tile edge: what a finished pixel still collects is below 1e-4 per channel."""
    g = _awkward_scene(40000, seed=27)
    W, H = 733, 411
    r, _ = _setup(R, g, W, H, eye=(1.0, -0.5, 9.0), mod=mod, sm=sm, debug=False)
    imgs = []
    for tile in (32, 16):
        r.tune(_gsr.GSR_TUNE_TILE, tile)
        imgs.append(r.draw().cpu().numpy().copy())
        s = r.stats()
        assert (s.tiles_x, s.tiles_y) == ((W + tile - 1) // tile, (H + tile - 1) // tile)
    if mod in (-2, -3, -4):
        assert np.array_equal(_bits(imgs[0]), _bits(imgs[1]))
    else:
        d = np.abs(imgs[0] - imgs[1])
        _report(f"[parity] tile 32 vs 16, mod {mod}: max|d| {d.max():.3e}, pixels that differ {int((d.max(axis=-1) > 0).sum())}")
        assert d.max() <= 2e-4


def test_fast_and_generic_depth_sort_agree_bitwise(R):
    """The frame's fast depth sort (digit counts from the preprocess kernel, a compacting first pass, three 9-bit
    passes over the rebased keys of the visible Gaussians) against the generic one (histogram kernel, four 8-bit
    passes over all N keys with the culled ones at the end): instance list, ranges and image bit-identical
    (ragged N: last span partial; a scene where most Gaussians are culled; per-handle knob)."""
    for n, eye in [(300_007, (0.0, 0.0, 10.0)), (200_003, (0.0, 0.0, 2.5)), (5_001, (3.0, 1.0, 6.0))]:
        g = _awkward_scene(n, seed=31)
        W, H = 1280, 720
        out = []
        for fast in (1, 0):
            r, _ = _setup(R, g, W, H, eye=eye, debug=False)
            r.tune(_gsr.GSR_TUNE_FUSED_DEPTH_HIST, fast)
            img = r.draw().cpu().numpy()
            out.append((_bits(img), r.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32),
                        r.debug_fetch(_gsr.GSR_DBG_RANGES, np.int32)))
            r2, _ = _setup(R, g, W, H, eye=eye, debug=False)   # another handle keeps its own (default) setting
            assert np.array_equal(_bits(r2.draw().cpu().numpy()), out[0][0])
        for a_, b_ in zip(out[0], out[1]):
            assert np.array_equal(a_, b_)
        assert len(out[0][1]) > 0


@pytest.mark.parametrize("n,eye_z", [(1_000_003, 60.0), (2_200_001, 60.0), (3_000_000, 60.0), (3_500_003, 60.0),
                                     (2_158_000, 10.0), (444 * 4864 + 5, 80.0), (444 * 7680 + 31, 80.0)])
def test_depth_sort_span_shapes_agree_with_the_generic_sort(R, n, eye_z):
    """The compacting depth sort cuts the visible keys into one span per resident CTA: one sub-tile of 27 (first
    pass) / 19 (other passes) keys per thread, a second sub-tile when more than 444 x 4864 Gaussians are visible, a second
    wave above 444 x 7680 (B200: 148 SMs x 3 CTAs).  With the camera far away every Gaussian is visible, so N picks the
    shape: ragged last warps, exactly full sub-tiles, the sub-tile and wave boundaries +- a few keys.  Instance list and
    ranges bit-identical to the generic sort (histogram kernel, four 8-bit passes over all N)."""
    g = scenes.synthetic_gaussians(n, seed=37)
    W, H = 640, 360
    out = []
    for fast in (1, 0):
        r, _ = _setup(R, g, W, H, eye=(0.0, 0.0, eye_z), mod=0, debug=False)
        r.tune(_gsr.GSR_TUNE_FUSED_DEPTH_HIST, fast)
        r.draw()
        st = r.stats()
        out.append((r.debug_fetch(_gsr.GSR_DBG_INST_VALS, np.uint32), r.debug_fetch(_gsr.GSR_DBG_RANGES, np.int32),
                    np.array([st.n_visible, st.n_instances])))
        del r
    assert out[0][2][0] > 0.6 * n or eye_z < 20.0
    _report(f"[parity] depth-sort shapes: N {n}, visible {int(out[0][2][0])}, instances {int(out[0][2][1])}")
    for a_, b_ in zip(out[0], out[1]):
        assert np.array_equal(a_, b_)


@pytest.mark.parametrize("W,H,tile", [(483, 271, 32), (484, 270, 16), (96, 64, 32), (50, 30, 32)])
def test_rgb8_from_the_blend_kernel_matches_the_pack_kernel(R, W, H, tile):
    """gsr_render_rgb8_host_async takes the RGB8 frame from the blend kernel's epilogue (aligned words out of
    shared memory for full tiles, byte stores for tiles cut by the image edge or widths that are not a multiple
    of four); the synchronous read-back converts the float frame with k_pack_rgb8.  Same bytes."""
    g = scenes.synthetic_gaussians(30000, seed=33)
    r, _ = _setup(R, g, W, H, debug=False)
    r.tune(_gsr.GSR_TUNE_TILE, tile)
    want = r.read_pixels_rgb8().copy()
    got = r.read_pixels_rgb8_async()
    r.sync_host()
    assert want.shape == (H, W, 3) and want.max() > 0
    assert np.array_equal(got, want)
    r.tune(_gsr.GSR_TUNE_FUSED_RGB8, 0)
    try:
        got2 = r.read_pixels_rgb8_async()
        r.sync_host()
        assert np.array_equal(got2, want)
    finally:
        r.tune(_gsr.GSR_TUNE_FUSED_RGB8, 1)
