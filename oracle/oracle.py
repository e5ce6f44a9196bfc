"""TEST INFRASTRUCTURE — not product code.

ctypes bindings for the two CPU checkers:

* ``libgsoracle.so``  — oracle/gs_oracle.c, the committed C restatement of the reference's render
  path (builds anywhere gcc is present; travels to the GPU box as a prebuilt .so too).
* ``_ref/libgsref.so`` — the reference's own shader source compiled as C++ (oracle/ref_build.py);
  can only be (re)built where /root/reference exists, but the built .so travels.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  Nothing under gaussiansplattingviewer_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libgsoracle.so")
REF_SO = os.path.join(HERE, "_ref", "libgsref.so")


class Camera(C.Structure):
    """GsoCamera / GsRefCamera (identical layout)."""
    _fields_ = [
        ("view", C.c_float * 16), ("proj", C.c_float * 16), ("hfovxy_focal", C.c_float * 3),
        ("cam_pos", C.c_float * 3), ("scale_modifier", C.c_float), ("render_mod", C.c_int), ("sh_dim", C.c_int),
    ]


VS_DTYPE = np.dtype([
    ("visible", "<f4"), ("ndc", "<f4", 3), ("half", "<f4", 2), ("conic", "<f4", 3), ("color", "<f4", 3),
    ("alpha", "<f4"), ("depth", "<f4"), ("pad", "<f4", 2),
])
SETUP_DTYPE = np.dtype([
    ("cx", "<f4"), ("cy", "<f4"), ("depth", "<f4"), ("ok", "<i4"), ("imin", "<i4"), ("imax", "<i4"),
    ("jmin", "<i4"), ("jmax", "<i4"),
])


def make_camera(view, proj, hfovxy_focal, cam_pos, scale_modifier=1.0, render_mod=3, sh_dim=48) -> Camera:
    cam = Camera()
    cam.view[:] = np.asarray(view, dtype=np.float32).reshape(16).tolist()
    cam.proj[:] = np.asarray(proj, dtype=np.float32).reshape(16).tolist()
    cam.hfovxy_focal[:] = np.asarray(hfovxy_focal, dtype=np.float32).reshape(3).tolist()
    cam.cam_pos[:] = np.asarray(cam_pos, dtype=np.float32).reshape(-1)[:3].tolist()
    cam.scale_modifier = float(scale_modifier)
    cam.render_mod = int(render_mod)
    cam.sh_dim = int(sh_dim)
    return cam


def build_oracle(force: bool = False) -> str:
    src = os.path.join(HERE, "gs_oracle.c")
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared", src, "-o",
               ORACLE_SO, "-lm"]
        subprocess.run(cmd, check=True)
    return ORACLE_SO


def build_ref(force: bool = False):
    """Builds oracle/_ref when the reference tree is present; returns the path or None."""
    if os.path.exists(REF_SO) and not force:
        return REF_SO
    from . import ref_build
    return ref_build.build()


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a, t=C.c_float):
    return a.ctypes.data_as(C.POINTER(t))


_oracle = None
_ref = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        build_oracle()
        L = C.CDLL(ORACLE_SO)
        L.gso_tile_instances.restype = C.c_longlong
        _oracle = L
    return _oracle


def ref_lib():
    """The reference-shader build, or None if it was never built and cannot be built here."""
    global _ref
    if _ref is None:
        so = build_ref()
        if so is None or not os.path.exists(so):
            return None
        _ref = C.CDLL(so)
    return _ref


# ---- C restatement -----------------------------------------------------------------------------------
def preprocess(gau, cam: Camera, W: int, H: int):
    """-> (vs[N] VS_DTYPE, setup[N] SETUP_DTYPE)"""
    n = len(gau)
    xyz, rot, scale, op, sh = _f32(gau.xyz), _f32(gau.rot), _f32(gau.scale), _f32(gau.opacity), _f32(gau.sh)
    vs = np.zeros(n, dtype=VS_DTYPE)
    st = np.zeros(n, dtype=SETUP_DTYPE)
    oracle_lib().gso_preprocess(_p(xyz), _p(rot), _p(scale), _p(op), _p(sh), C.c_int(n), C.byref(cam), C.c_int(W),
                                C.c_int(H), vs.ctypes.data_as(C.c_void_p), st.ctypes.data_as(C.c_void_p))
    return vs, st


def sort_depth(xyz, view):
    """Stable restatement of _sort_gaussian_cpu: (index (N,1) int32 far->near, depth (N,) f32)."""
    xyz = _f32(xyz)
    n = len(xyz)
    view = _f32(view).reshape(16)
    idx = np.zeros(n, dtype=np.int32)
    depth = np.zeros(n, dtype=np.float32)
    oracle_lib().gso_sort_depth(_p(xyz), C.c_int(n), _p(view), _p(idx, C.c_int), _p(depth))
    return idx.reshape(-1, 1), depth


def sort_pairs(keys: np.ndarray, vals: np.ndarray):
    keys = np.ascontiguousarray(keys).copy()
    vals = np.ascontiguousarray(vals, dtype=np.uint32).copy()
    if keys.dtype == np.uint64:
        oracle_lib().gso_sort_pairs_u64(_p(keys, C.c_uint64), _p(vals, C.c_uint32), C.c_longlong(len(keys)))
    else:
        keys = keys.astype(np.uint32)
        oracle_lib().gso_sort_pairs_u32(_p(keys, C.c_uint32), _p(vals, C.c_uint32), C.c_longlong(len(keys)))
    return keys, vals


def visible_order(setup):
    n = len(setup)
    out = np.zeros(max(n, 1), dtype=np.int32)
    m = oracle_lib().gso_visible_order(setup.ctypes.data_as(C.c_void_p), C.c_int(n), _p(out, C.c_int))
    return out[:m].copy()


def tile_instances(setup, W, H, tile=16, vs=None, cull=False):
    """-> (keys u64 sorted, vals u32, ranges (tiles, 2) int32).  cull: apply the oracle/gs_tilecull.h rule
    (needs the vertex-stage outputs `vs`)."""
    L = oracle_lib()
    n = len(setup)
    vsp = vs.ctypes.data_as(C.c_void_p) if vs is not None else None
    cl = C.c_int(1 if (cull and vs is not None) else 0)
    total = L.gso_tile_instances(setup.ctypes.data_as(C.c_void_p), vsp, cl, C.c_int(n), C.c_int(W), C.c_int(H),
                                 C.c_int(tile), None, None, C.c_longlong(0), None)
    keys = np.zeros(max(total, 1), dtype=np.uint64)
    vals = np.zeros(max(total, 1), dtype=np.uint32)
    tiles = ((W + tile - 1) // tile) * ((H + tile - 1) // tile)
    ranges = np.zeros((tiles, 2), dtype=np.int32)
    got = L.gso_tile_instances(setup.ctypes.data_as(C.c_void_p), vsp, cl, C.c_int(n), C.c_int(W), C.c_int(H), C.c_int(tile),
                               _p(keys, C.c_uint64), _p(vals, C.c_uint32), C.c_longlong(total), _p(ranges, C.c_int))
    assert got == total
    return keys[:total], vals[:total], ranges


def render(gau, cam: Camera, W: int, H: int, clamp_unorm: bool = True, want_fragile: bool = False):
    """Whole frame -> (rgba (H,W,4) f32 top-down, fragile (H,W) u8 or None, n_drawn)."""
    n = len(gau)
    xyz, rot, scale, op, sh = _f32(gau.xyz), _f32(gau.rot), _f32(gau.scale), _f32(gau.opacity), _f32(gau.sh)
    img = np.zeros((H, W, 4), dtype=np.float32)
    frag = np.zeros((H, W), dtype=np.uint8) if want_fragile else None
    m = oracle_lib().gso_render(_p(xyz), _p(rot), _p(scale), _p(op), _p(sh), C.c_int(n), C.byref(cam), C.c_int(W),
                                C.c_int(H), C.c_int(1 if clamp_unorm else 0), _p(img),
                                frag.ctypes.data_as(C.c_void_p) if want_fragile else None)
    return img, frag, m


def num_threads() -> int:
    return int(oracle_lib().gso_num_threads())


def set_num_threads(n: int) -> None:
    """OpenMP threads of the calling thread's next parallel regions (both oracle libraries share one libgomp)."""
    oracle_lib().gso_set_num_threads(int(n))


# ---- reference shaders compiled as C++ ---------------------------------------------------------------
def ref_vertex(gau, cam: Camera):
    L = ref_lib()
    if L is None:
        raise RuntimeError("oracle/_ref/libgsref.so is not available")
    flat = _f32(gau.flat())
    n = len(gau)
    out = np.zeros(n, dtype=VS_DTYPE)
    L.gsref_vertex(_p(flat), C.c_int(n), C.byref(cam), out.ctypes.data_as(C.c_void_p))
    return out


def ref_render(gau, order, cam: Camera, W: int, H: int, clamp_unorm: bool = True):
    L = ref_lib()
    if L is None:
        raise RuntimeError("oracle/_ref/libgsref.so is not available")
    flat = _f32(gau.flat())
    n = len(gau)
    order = np.ascontiguousarray(np.asarray(order).reshape(-1), dtype=np.int32)
    assert len(order) == n
    img = np.zeros((H, W, 4), dtype=np.float32)
    L.gsref_render(_p(flat), C.c_int(n), _p(order, C.c_int), C.byref(cam), C.c_int(W), C.c_int(H),
                   C.c_int(1 if clamp_unorm else 0), _p(img))
    return img


# ---- official-rasterizer semantics (parity unpinned; see the header of that section in gs_oracle.c) ---------
def official_matrices(view, proj):
    """What CUDARenderer.update_camera_pose builds (/root/reference/renderer_cuda.py:181-194): the view with rows
    0 and 2 negated and P @ view, in binary32 with left-to-right sums (the arithmetic of gsr_api.cu)."""
    vo = np.array(view, dtype=np.float32, copy=True).reshape(4, 4)
    vo[[0, 2], :] = -vo[[0, 2], :]
    P = np.asarray(proj, dtype=np.float32).reshape(4, 4)
    po = np.zeros((4, 4), dtype=np.float32)
    for r in range(4):
        for c in range(4):
            acc = np.float32(P[r, 0] * vo[0, c])
            for k in range(1, 4):
                acc = np.float32(acc + np.float32(P[r, k] * vo[k, c]))
            po[r, c] = acc
    return vo, po


def official_camera(view, proj, hfovxy_focal, cam_pos, scale_modifier=1.0, sh_dim=48) -> Camera:
    vo, po = official_matrices(view, proj)
    return make_camera(vo, po, hfovxy_focal, cam_pos, scale_modifier, 3, sh_dim)


def official_preprocess(gau, cam: Camera, W: int, H: int):
    n = len(gau)
    xyz, rot, scale, op, sh = _f32(gau.xyz), _f32(gau.rot), _f32(gau.scale), _f32(gau.opacity), _f32(gau.sh)
    vs = np.zeros(n, dtype=VS_DTYPE)
    st = np.zeros(n, dtype=SETUP_DTYPE)
    oracle_lib().gso_official_preprocess(_p(xyz), _p(rot), _p(scale), _p(op), _p(sh), C.c_int(n), C.byref(cam),
                                         C.c_int(W), C.c_int(H), vs.ctypes.data_as(C.c_void_p),
                                         st.ctypes.data_as(C.c_void_p))
    return vs, st


def official_render(gau, cam: Camera, W: int, H: int, want_fragile: bool = False):
    """-> (rgba (H,W,4) f32 top-down with alpha 1, fragile (H,W) u8 or None, n_drawn)"""
    n = len(gau)
    xyz, rot, scale, op, sh = _f32(gau.xyz), _f32(gau.rot), _f32(gau.scale), _f32(gau.opacity), _f32(gau.sh)
    img = np.zeros((H, W, 4), dtype=np.float32)
    frag = np.zeros((H, W), dtype=np.uint8) if want_fragile else None
    m = oracle_lib().gso_official_render(_p(xyz), _p(rot), _p(scale), _p(op), _p(sh), C.c_int(n), C.byref(cam),
                                         C.c_int(W), C.c_int(H), _p(img),
                                         frag.ctypes.data_as(C.c_void_p) if want_fragile else None)
    return img, frag, m
