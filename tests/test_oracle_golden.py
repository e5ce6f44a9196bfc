"""The oracle (oracle/gs_oracle.c) against the reference: golden vectors generated from the
reference's own code (tests/golden/make_golden.py) and, where oracle/_ref/libgsref.so is present,
the reference shaders run live.  CPU only."""
import numpy as np
import pytest

from gaussiansplattingviewer_b200 import scenes, util, util_gau
from gaussiansplattingviewer_b200.util_gau import GaussianData
from oracle import oracle as O

VS_FIELDS = ["ndc", "half", "conic", "color", "alpha"]


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _vs_as_rows(vs):
    return vs.view(np.float32).reshape(-1, 16)[:, :13]


def test_vertex_stage_matches_reference_shader_goldens(golden):
    z = golden("vertex_ref.npz")
    g = GaussianData(z["xyz"], z["rot"], z["scale"], z["opacity"], z["sh"])
    W, H = int(z["W"]), int(z["H"])
    keys = [k for k in z.files if k.startswith("vs_mod")]
    assert len(keys) == 10
    for k in keys:
        mod = int(k.split("_")[1][3:])
        sm = float(k.split("_sm")[1])
        cam = O.make_camera(z["view"], z["proj"], z["hff"], z["campos"], sm, mod, g.sh_dim)
        vs, _ = O.preprocess(g, cam, W, H)
        got, want = _vs_as_rows(vs), z[k]
        assert np.array_equal(got[:, 0], want[:, 0]), k  # visibility
        vis = want[:, 0] > 0
        assert vis.sum() > 1000
        assert np.array_equal(_bits(got[vis]), _bits(want[vis])), k  # bit-exact


def test_config1_images_match_reference_shader_goldens(golden):
    z = golden("naive_images_ref.npz")
    g, _, _ = util_gau.naive_gaussian()
    for mod in [3, 0, -1, -2, -3, -4]:
        cam = O.make_camera(z["view"], z["proj"], z["hff"], z["campos"], 1.0, mod, g.sh_dim)
        img, _, drawn = O.render(g, cam, 1280, 720, clamp_unorm=(mod != -1))
        assert drawn == 4
        want = z[f"img_mod{mod}"]
        assert want[..., :3].max() > 0.5 or mod == -1
        assert np.array_equal(_bits(img), _bits(want)), mod


def test_synthetic_images_match_reference_shader_goldens(golden):
    z = golden("synth_images_ref.npz")
    g = scenes.synthetic_gaussians(int(z["n"]), seed=int(z["seed"]))
    W, H = int(z["W"]), int(z["H"])
    for mod in [3, -1, -3, -4]:
        cam = O.make_camera(z["view"], z["proj"], z["hff"], z["campos"], 1.0, mod, g.sh_dim)
        img, _, _ = O.render(g, cam, W, H, clamp_unorm=(mod != -1))
        want = z[f"img_mod{mod}"]
        # same arithmetic, same order (no depth ties in this scene): identical
        assert np.array_equal(_bits(img), _bits(want)), (mod, np.abs(img - want).max())


def test_depth_sort_matches_reference_sort_goldens(golden):
    z = golden("sort_ref.npz")
    # (a) generic floats: depths distinct -> the index order is unique up to float rounding of the
    # reference's numpy matmul; require the reference order to be a valid ascending order of OUR depths
    idx, depth = O.sort_depth(z["xyz_a"], z["view_a"])
    assert idx.dtype == np.int32 and idx.shape == z["index_a"].shape
    ref = z["index_a"][:, 0]
    assert np.array_equal(np.sort(ref), np.arange(len(ref)))
    mism = idx[:, 0] != ref
    # any disagreement must be between depths that differ by rounding only
    if mism.any():
        assert np.abs(depth[idx[mism, 0]] - depth[ref[mism]]).max() < 1e-5
    assert mism.mean() < 0.01
    # (b) exactly representable arithmetic, many ties: sorted depth sequence identical, and our order is
    # the STABLE one (ascending index inside every tie group)
    idx, depth = O.sort_depth(z["xyz_b"], z["view_b"])
    assert np.array_equal(depth, z["depth_b"])
    assert np.array_equal(depth[idx[:, 0]], z["depth_b"][z["index_b"][:, 0]])
    want = np.lexsort((np.arange(len(depth)), depth))
    assert np.array_equal(idx[:, 0], want)


def test_sort_special_values():
    xyz = np.zeros((8, 3), dtype=np.float32)
    xyz[:, 2] = [np.nan, -0.0, 0.0, np.inf, -np.inf, 1e-40, -1e-40, -np.nan]
    view = np.eye(4, dtype=np.float32)
    idx, depth = O.sort_depth(xyz, view)
    # numpy semantics: NaN last, -0 == +0 (stable), denormals ordered
    want = np.argsort(xyz[:, 2], kind="stable")
    assert np.array_equal(idx[:, 0], want)


def test_tile_instances_are_lexsorted():
    g = scenes.synthetic_gaussians(3000, seed=3)
    W, H = 200, 120
    cam, pose = scenes.default_view(W, H)
    c = O.make_camera(pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal(), pose["camera_position"])
    _, st = O.preprocess(g, c, W, H)
    keys, vals, ranges = O.tile_instances(st, W, H, 16)
    tiles_x = (W + 15) // 16
    exp_t, exp_d, exp_g = [], [], []
    for gi in np.nonzero(st["ok"])[0]:
        s = st[gi]
        for ty in range(s["jmin"] // 16, s["jmax"] // 16 + 1):
            for tx in range(s["imin"] // 16, s["imax"] // 16 + 1):
                exp_t.append(ty * tiles_x + tx)
                exp_d.append(np.float32(s["depth"]).view(np.uint32))
                exp_g.append(gi)
    exp_t, exp_d, exp_g = np.array(exp_t), np.array(exp_d, dtype=np.uint64), np.array(exp_g)
    o = np.lexsort((exp_g, exp_d, exp_t))
    assert len(keys) == len(o) > 3000
    assert np.array_equal(vals, exp_g[o].astype(np.uint32))
    assert np.array_equal(keys, (exp_t[o].astype(np.uint64) << np.uint64(32)) | exp_d[o])
    for t in range(len(ranges)):
        a, b = ranges[t]
        assert np.all((keys[a:b] >> np.uint64(32)) == t)
    assert (ranges[:, 1] - ranges[:, 0]).sum() == len(keys)


def test_known_answer_single_isotropic_splat():
    # isotropic splat on the optical axis: a = c = (focal*s/z)^2 + 0.3, b = 0,
    # half extent 3*sqrt(a), alpha at the centre pixel = min(0.99, opacity * exp(-small))
    W, H = 1281, 721  # odd so a pixel centre sits exactly on the axis
    cam = util.Camera(H, W)
    V, P, hff = cam.get_view_matrix(False), cam.get_project_matrix(), cam.get_htanfovxy_focal()
    s, z, op = 0.05, 3.0, 0.6
    g = GaussianData(np.zeros((1, 3), np.float32), np.array([[1, 0, 0, 0]], np.float32),
                     np.full((1, 3), s, np.float32), np.full((1, 1), op, np.float32),
                     np.array([[1.0, 0.5, -0.2]], np.float32))
    c = O.make_camera(V, P, hff, cam.position, 1.0, 0, 3)
    vs, st = O.preprocess(g, c, W, H)
    a = (hff[2] * s / z) ** 2 + 0.3
    assert abs(vs["half"][0, 0] - 3 * np.sqrt(a)) < 1e-3
    assert abs(vs["conic"][0, 0] - 1 / a) < 1e-7 and abs(vs["conic"][0, 1]) < 1e-9
    assert st["ok"][0] == 1 and abs(st["cx"][0] - W / 2) < 1e-3 and abs(st["cy"][0] - H / 2) < 1e-3
    img, _, _ = O.render(g, c, W, H)
    centre = img[H // 2, W // 2]
    col = np.clip(0.28209479177387814 * np.array([1.0, 0.5, -0.2]) + 0.5, 0, 1)
    assert abs(centre[3] - op) < 1e-5
    assert np.allclose(centre[:3], op * col, atol=1e-5)
    # outside the 3-sigma box nothing is drawn
    r = int(np.ceil(3 * np.sqrt(a))) + 2
    assert img[H // 2, W // 2 + r, 3] == 0 and img[H // 2 + r, W // 2, 3] == 0


@pytest.mark.parametrize("mod,sm", [(3, 1.0), (-1, 1.0), (-3, 1.7), (-4, 0.5), (-2, 1.0), (1, 1.0)])
def test_oracle_against_live_reference_shaders(mod, sm):
    if O.ref_lib() is None:
        pytest.skip("oracle/_ref/libgsref.so not built (needs /root/reference)")
    g = scenes.synthetic_gaussians(6000, seed=100 + mod)
    W, H = 384, 216
    cam, pose = scenes.default_view(W, H, eye=(2.0, 1.0, 7.0))
    V, P, hff = pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal()
    c = O.make_camera(V, P, hff, pose["camera_position"], sm, mod, g.sh_dim)
    vs, st = O.preprocess(g, c, W, H)
    rv = O.ref_vertex(g, c)
    assert np.array_equal(vs["visible"], rv["visible"])
    vis = rv["visible"] > 0
    for f in VS_FIELDS:
        assert np.array_equal(_bits(vs[f][vis]), _bits(rv[f][vis])), f
    img, _, _ = O.render(g, c, W, H, clamp_unorm=(mod != -1))
    front_to_back = O.visible_order(st)
    rest = np.setdiff1d(np.arange(len(g)), front_to_back)
    order = np.concatenate([rest, front_to_back[::-1]]).astype(np.int32)
    rimg = O.ref_render(g, order, c, W, H, clamp_unorm=(mod != -1))
    assert np.array_equal(_bits(img), _bits(rimg))
