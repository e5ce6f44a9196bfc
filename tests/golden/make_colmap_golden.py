"""Generates tests/golden/colmap_ref.npz FROM THE REFERENCE's own pose code, in the build container.

Run:  python tests/golden/make_colmap_golden.py        (needs /root/reference; not run on the GPU box)

/root/reference/main.py cannot be imported (glfw, imgui, cv2, tkinter, torchmetrics ... at module scope), so
the functions of the pose path are cut out of its syntax tree and executed unmodified:

    getWorld2View2 (:34), qvec2rotmat (:102), quaternion_to_rotation_matrix (:167), apply_rotation (:185),
    create_look_at_from_colmap (:197), load_camera_positions (:275-407)

`quaternion_to_rotation_matrix` / `create_look_at_from_colmap` are pure NumPy: their outputs are pinned bit for
bit.  `load_camera_positions` additionally calls PyGLM (`glm.lookAtRH`, `glm.mat4`, `*`, `glm.inverse`), which is
not installed: a small NumPy-backed `glm` shim (below; GLM's published float32 formulas, column-major indexing)
stands in, so the view matrices are pinned to the REFERENCE'S CALL SEQUENCE run on that shim, not to a PyGLM
binary — the test compares them to 1 ulp / 1e-6.
"""
from __future__ import annotations

import ast
import os
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_MAIN = "/root/reference/main.py"
WANT = ["getWorld2View2", "qvec2rotmat", "quaternion_to_rotation_matrix", "apply_rotation", "normalize_vector",
        "create_look_at_from_colmap", "load_camera_positions"]

# the three pose rows of tests/test_colmap_poses.py (IMAGE_ID QW QX QY QZ TX TY TZ CAMERA_ID NAME)
ROWS = [
    "1 0.851773 0.0165051 0.503764 -0.142941 -0.737434 1.02973 3.74354 1 a.jpg",
    "2 1 0 0 0 0.5 -1.5 2.0 1 b.jpg",
    "3 0.5 -0.5 0.5 0.5 3.0 0.25 -4.0 1 c.jpg",
]

f32 = np.float32


class Mat4:
    """glm.mat4 stand-in: column-major indexing m[col, row] / m[col] over a maths-layout float32 array."""

    def __init__(self, a):
        self.a = np.array(a, dtype=np.float32).reshape(4, 4)

    def __getitem__(self, k):
        if isinstance(k, tuple):
            return self.a[k[1], k[0]]
        return self.a[:, k].copy()  # column k

    def __setitem__(self, k, v):
        self.a[k[1], k[0]] = f32(v)

    def __mul__(self, o):
        # glm: column c of A*B = A[0]*B[c][0] + A[1]*B[c][1] + A[2]*B[c][2] + A[3]*B[c][3], float32, left to right
        out = np.zeros((4, 4), dtype=np.float32)
        for c in range(4):
            acc = self.a[:, 0] * o.a[0, c]
            for k in range(1, 4):
                acc = (acc + self.a[:, k] * o.a[k, c]).astype(np.float32)
            out[:, c] = acc
        return Mat4(out)

    def __array__(self, dtype=None, copy=None):
        return self.a.astype(dtype) if dtype is not None else self.a.copy()


def _vec3(*a):
    if len(a) == 1:
        return np.array(np.asarray(a[0], dtype=np.float64).reshape(-1)[:3], dtype=np.float32)
    return np.array(a, dtype=np.float32)


def _normalize(v):
    # glm::normalize: v * inversesqrt(dot(v, v)) in float32
    d = f32(f32(f32(v[0] * v[0]) + f32(v[1] * v[1])) + f32(v[2] * v[2]))
    return (v * f32(f32(1.0) / np.sqrt(d, dtype=np.float32))).astype(np.float32)


def _cross(x, y):
    return np.array([f32(x[1] * y[2]) - f32(y[1] * x[2]), f32(x[2] * y[0]) - f32(y[2] * x[0]),
                     f32(x[0] * y[1]) - f32(y[0] * x[1])], dtype=np.float32)


def _dot(x, y):
    return f32(f32(f32(x[0] * y[0]) + f32(x[1] * y[1])) + f32(x[2] * y[2]))


def _look_at_rh(eye, center, up):
    # glm::lookAtRH (matrix_transform.inl)
    f = _normalize((center - eye).astype(np.float32))
    s = _normalize(_cross(f, up))
    u = _cross(s, f)
    m = np.eye(4, dtype=np.float32)
    m[0, :3] = s
    m[1, :3] = u
    m[2, :3] = -f
    m[0, 3] = -_dot(s, eye)
    m[1, 3] = -_dot(u, eye)
    m[2, 3] = _dot(f, eye)
    return Mat4(m)


def make_glm():
    glm = types.ModuleType("glm")
    glm.vec3 = _vec3
    glm.vec4 = lambda v: np.array(v, dtype=np.float32)
    glm.mat4 = lambda *a: Mat4(np.eye(4) * (a[0] if a else 1.0))
    glm.quat = lambda *a: tuple(a)
    glm.mat4_cast = lambda q: None  # result unused by the reference (:329)
    glm.lookAtRH = _look_at_rh
    glm.lookAt = _look_at_rh
    glm.inverse = lambda m: Mat4(np.linalg.inv(m.a.astype(np.float64)))
    glm.normalize = _normalize
    return glm


def load_reference_functions():
    with open(REF_MAIN) as fh:
        tree = ast.parse(fh.read())
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in WANT]
    assert sorted(n.name for n in body) == sorted(WANT), [n.name for n in body]
    mod = ast.Module(body=body, type_ignores=[])
    glm = make_glm()
    ns = {"np": np, "glm": glm, "debug_vector": False, "new_camera_pos": None,
          "up_vector": glm.vec3(0, -1, 0), "front_vector": glm.vec3(0, 0, 1)}  # module globals, main.py:83,160-161
    exec(compile(mod, REF_MAIN, "exec"), ns)
    return ns


def main():
    ns = load_reference_functions()
    out = {"rows": np.array(ROWS)}
    for i, row in enumerate(ROWS):
        e = row.split()
        qw, qx, qy, qz = (float(e[k]) for k in (1, 2, 3, 4))
        x, y, z = (float(e[k]) for k in (5, 6, 7))
        out[f"rotmat_{i}"] = ns["quaternion_to_rotation_matrix"](qw, qx, qy, qz)
        pos, center, up = ns["create_look_at_from_colmap"](x, y, z, qw, qx, qy, qz)
        out[f"pos_{i}"], out[f"center_{i}"], out[f"up_{i}"] = pos, center, up
        pl, pr = ns["load_camera_positions"](e)
        out[f"view_left_{i}"] = np.array(pl["camera_view"], dtype=np.float32)
        out[f"view_right_{i}"] = np.array(pr["camera_view"], dtype=np.float32)
        out[f"campos_left_{i}"] = np.asarray(pl["camera_position"], dtype=np.float64)
        out[f"campos_right_{i}"] = np.asarray(pr["camera_position"], dtype=np.float32)[:3]
        out[f"front_{i}"] = np.asarray(pl["camera_front"], dtype=np.float32)
        out[f"up_pose_{i}"] = np.asarray(pl["camera_up"], dtype=np.float64)
    np.savez(os.path.join(HERE, "colmap_ref.npz"), **out)
    print("colmap_ref.npz", os.path.getsize(os.path.join(HERE, "colmap_ref.npz")))


if __name__ == "__main__":
    main()
