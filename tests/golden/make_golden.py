"""Generates the golden fixtures under tests/golden/ FROM THE REFERENCE, in the build container.

Run:  python tests/golden/make_golden.py          (needs /root/reference; not run on the GPU box)

Two sources, both the reference's own code:

1. Python functions imported unmodified from /root/reference with stub modules standing in for the
   packages that are not installed (OpenGL, glm, plyfile, open3d):
     renderer_ogl._sort_gaussian_cpu      -> sort_ref.npz
     util_gau.GaussianData / naive_gaussian -> naive_ref.npz
     util.Camera.get_project_matrix / get_htanfovxy_focal -> camera_ref.npz
     util_gau.load_ply (fed by a stub PlyData that serves columns read by our own container
     parser; what is pinned is the reference's column selection, re-ordering and activations)
                                          -> ply_ref.npz (+ tiny.ply)
   camera_data.csv (data, 18 poses)       -> camera_data.npy
2. The reference's shader source compiled as C++ (oracle/ref_build.py -> oracle/_ref/libgsref.so):
     vertex-stage outputs for a seeded scene, every render mode        -> vertex_ref.npz
     config-1 images (naive_gaussian, 1280x720, eye (0,0,3)), 6 modes  -> naive_images_ref.npz
     a seeded 4k-splat scene at 320x180, 4 modes                       -> synth_images_ref.npz
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

from gaussiansplattingviewer_b200 import scenes, util as my_util, util_gau as my_gau  # noqa: E402
from oracle import oracle as O  # noqa: E402


def _install_stubs():
    """Stand-ins for modules the reference imports at module scope but that the functions we call
    never touch (GL) or that only provide a container (plyfile)."""
    gl = types.ModuleType("OpenGL")
    gl_gl = types.ModuleType("OpenGL.GL")
    gl_sh = types.ModuleType("OpenGL.GL.shaders")
    gl.GL = gl_gl
    gl_gl.shaders = gl_sh
    sys.modules.update({"OpenGL": gl, "OpenGL.GL": gl_gl, "OpenGL.GL.shaders": gl_sh})

    glm = types.ModuleType("glm")
    glm.vec3 = lambda *a: np.array(a, dtype=np.float32)
    glm.radians = np.radians
    glm.angleAxis = lambda ang, ax: None
    glm.perspective = lambda *a: None  # result is discarded by the reference (util.py:94-105)
    sys.modules["glm"] = glm

    o3d = types.ModuleType("open3d")
    sys.modules["open3d"] = o3d

    ply = types.ModuleType("plyfile")

    class _Prop:
        def __init__(self, name):
            self.name = name

    class _Elem:
        def __init__(self, arr):
            self._a = arr
            self.properties = [_Prop(n) for n in arr.dtype.names]

        def __getitem__(self, k):
            return self._a[k]

    class PlyData:
        def __init__(self, arr):
            self.elements = [_Elem(arr)]

        @staticmethod
        def read(path):
            return PlyData(my_gau.read_ply_vertices(path))

    ply.PlyData = PlyData
    sys.modules["plyfile"] = ply


def main():
    _install_stubs()
    sys.path.insert(0, REF)
    import util_gau as ref_gau  # noqa
    import util as ref_util  # noqa
    import renderer_ogl as ref_ogl  # noqa

    # ---- 1a. naive_gaussian / flat ----
    g, a, b = ref_gau.naive_gaussian()
    np.savez(os.path.join(HERE, "naive_ref.npz"), xyz=g.xyz, rot=g.rot, scale=g.scale, opacity=g.opacity, sh=g.sh,
             flat=g.flat(), length=len(g), sh_dim=g.sh_dim, ret1=a, ret2=b)

    # ---- 1b. camera intrinsics ----
    cams = {}
    for (h, w) in [(720, 1280), (1080, 1920), (2160, 3840), (522, 1160)]:
        c = ref_util.Camera(h, w)
        cams[f"proj_{h}x{w}"] = c.get_project_matrix()
        cams[f"hff_{h}x{w}"] = np.array(c.get_htanfovxy_focal(), dtype=np.float64)
        cams[f"focal_{h}x{w}"] = np.float64(c.get_focal())
    c = ref_util.Camera(720, 1280)
    cams["fovy"] = np.float64(c.fovy)
    cams["position"] = c.position
    cams["target"] = c.target
    cams["up"] = c.up
    np.savez(os.path.join(HERE, "camera_ref.npz"), **cams)

    # ---- 1c. depth sort (reference _sort_gaussian_cpu) ----
    rng = np.random.default_rng(7)
    out = {}
    # (i) generic floats, distinct depths with overwhelming probability
    xyz = rng.normal(0, 3, size=(5000, 3)).astype(np.float32)
    _, pose = scenes.default_view(640, 360, eye=(1.0, 2.0, 9.0))
    view = pose["camera_view"].astype(np.float32)
    idx = ref_ogl._sort_gaussian_cpu(types.SimpleNamespace(xyz=xyz), view)
    out.update(xyz_a=xyz, view_a=view, index_a=idx)
    # (ii) exactly representable arithmetic (small integers / 8, axis-aligned view): no rounding anywhere
    xyz = (rng.integers(-64, 64, size=(4096, 3)) / 8.0).astype(np.float32)
    view = np.eye(4, dtype=np.float32)
    view[2, 3] = -12.0
    view[0, 0] = -1.0
    view[1, 1] = -1.0
    idx = ref_ogl._sort_gaussian_cpu(types.SimpleNamespace(xyz=xyz), view)
    depth = xyz[:, 2] + view[2, 3]
    out.update(xyz_b=xyz, view_b=view, index_b=idx, depth_b=depth)
    np.savez(os.path.join(HERE, "sort_ref.npz"), **out)

    # ---- 1d. load_ply ----
    n = 257
    xyz = rng.normal(0, 1.5, size=(n, 3)).astype(np.float32)
    f_dc = rng.normal(0, 1, size=(n, 3)).astype(np.float32)
    f_rest = rng.normal(0, 0.2, size=(n, 45)).astype(np.float32)
    op = rng.normal(0, 2, size=(n,)).astype(np.float32)
    ls = rng.normal(np.log(0.03), 0.7, size=(n, 3)).astype(np.float32)
    rot = rng.normal(0, 1, size=(n, 4)).astype(np.float32)
    ply_path = os.path.join(HERE, "tiny.ply")
    my_gau.save_ply(ply_path, xyz, f_dc, f_rest, op, ls, rot)
    gr, bbox, center = ref_gau.load_ply(ply_path)
    np.savez(os.path.join(HERE, "ply_ref.npz"), xyz=gr.xyz, rot=gr.rot, scale=gr.scale, opacity=gr.opacity, sh=gr.sh,
             bbox=bbox, center=center)

    # ---- 1e. camera_data.csv ----
    rows = np.loadtxt(os.path.join(REF, "camera_data.csv"), delimiter=",", dtype=np.float64)
    np.save(os.path.join(HERE, "camera_data.npy"), rows)

    # ---- 2. reference shaders compiled as C++ ----
    assert O.build_ref(force=True) is not None
    g = scenes.synthetic_gaussians(2048, seed=11)
    # push a few Gaussians into awkward places: behind the camera, on the guard band, huge, tiny
    g.xyz[:8] = np.array([[0, 0, 12], [0, 0, 9.95], [7.0, 0, 0], [0, 4.0, 0], [0, 0, -95], [0, 0, 9.89], [0.1, 0.1, 5],
                          [-6.9, -3.9, 0]], dtype=np.float32)
    g.scale[8:12] = np.array([[3, 3, 3], [1e-4, 1e-4, 1e-4], [2.0, 1e-3, 1e-3], [0.5, 0.5, 1e-3]], dtype=np.float32)
    W, H = 640, 360
    cam, pose = scenes.default_view(W, H)
    V, P, hff = pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal()
    vout = {"xyz": g.xyz, "rot": g.rot, "scale": g.scale, "opacity": g.opacity, "sh": g.sh, "view": V, "proj": P,
            "hff": np.array(hff, dtype=np.float32), "campos": pose["camera_position"], "W": W, "H": H}
    for mod, sm in [(3, 1.0), (2, 1.0), (1, 1.0), (0, 1.0), (-1, 1.0), (-2, 1.0), (-3, 1.0), (-4, 1.0), (3, 0.37), (3, 4.0)]:
        c = O.make_camera(V, P, hff, pose["camera_position"], sm, mod, g.sh_dim)
        vs = O.ref_vertex(g, c)
        vout[f"vs_mod{mod}_sm{sm}"] = vs.view(np.float32).reshape(-1, 16)[:, :13]
    np.savez_compressed(os.path.join(HERE, "vertex_ref.npz"), **vout)

    # config 1: naive_gaussian at 1280x720, non-arcball camera lookAt((0,0,3), 0, (0,-1,0))
    g, _, _ = my_gau.naive_gaussian()
    W, H = 1280, 720
    cam = my_util.Camera(H, W)
    V, P, hff = cam.get_view_matrix(False), cam.get_project_matrix(), cam.get_htanfovxy_focal()
    iout = {"view": V, "proj": P, "hff": np.array(hff, dtype=np.float32), "campos": cam.position}
    order = ref_ogl._sort_gaussian_cpu(g, V)
    iout["order"] = order
    for mod in [3, 0, -1, -2, -3, -4]:
        c = O.make_camera(V, P, hff, cam.position, 1.0, mod, g.sh_dim)
        img = O.ref_render(g, order, c, W, H, clamp_unorm=(mod != -1))
        iout[f"img_mod{mod}"] = img
    np.savez_compressed(os.path.join(HERE, "naive_images_ref.npz"), **iout)

    # seeded synthetic scene, small image
    g = scenes.synthetic_gaussians(4096, seed=5)
    W, H = 320, 180
    cam, pose = scenes.default_view(W, H, eye=(0.5, -1.0, 9.0))
    V, P, hff = pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal()
    sout = {"seed": 5, "n": 4096, "view": V, "proj": P, "hff": np.array(hff, dtype=np.float32),
            "campos": pose["camera_position"], "W": W, "H": H}
    order = ref_ogl._sort_gaussian_cpu(g, V)  # the reference's own (unstable on ties) far->near order
    sout["order"] = order
    for mod in [3, -1, -3, -4]:
        c = O.make_camera(V, P, hff, pose["camera_position"], 1.0, mod, g.sh_dim)
        sout[f"img_mod{mod}"] = O.ref_render(g, order, c, W, H, clamp_unorm=(mod != -1)).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "synth_images_ref.npz"), **sout)
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
