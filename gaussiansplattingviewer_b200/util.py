"""Camera mathematics of the viewer, without PyGLM / PyOpenGL.

Mirrors the math part of the reference ``Camera`` (/root/reference/util.py:10-116): same attribute
names and defaults, same ``get_view_matrix`` / ``get_project_matrix`` / ``get_htanfovxy_focal``
signatures and return layouts.  Mouse / keyboard handling (util.py:118-205) and the GL helper
functions (util.py:214-430) belong to the window and are out of scope (no graphics stack).

All matrices are NumPy arrays in *maths layout* (``M[row, col]``, translation in ``M[:3, 3]``),
which is what the reference hands to its sort (``view_mat[:3, 3]``, renderer_ogl.py:14) and
transposes before a GL upload (util.py:345).
"""
from __future__ import annotations

import math

import numpy as np


def look_at(eye, center, up) -> np.ndarray:
    """Right-handed look-at, the matrix ``glm.lookAt`` / ``glm.lookAtRH`` builds (float32 arithmetic
    like glm's vec3).  View space looks down -z.  Used where the reference calls
    ``np.array(glm.lookAt(...))`` (/root/reference/util.py:64-70)."""
    eye = np.asarray(eye, dtype=np.float32).reshape(3)
    center = np.asarray(center, dtype=np.float32).reshape(3)
    up = np.asarray(up, dtype=np.float32).reshape(3)
    f = center - eye
    f = f / np.float32(np.sqrt(np.dot(f, f)))
    s = np.cross(f, up).astype(np.float32)
    s = s / np.float32(np.sqrt(np.dot(s, s)))
    u = np.cross(s, f).astype(np.float32)
    m = np.eye(4, dtype=np.float32)
    m[0, :3] = s
    m[1, :3] = u
    m[2, :3] = -f
    m[0, 3] = -np.dot(s, eye)
    m[1, 3] = -np.dot(u, eye)
    m[2, 3] = np.dot(f, eye)
    return m


class Camera:
    """Reference: /root/reference/util.py:10-116 (state and math only)."""

    def __init__(self, h, w):
        self.znear = 0.1
        self.zfar = 100
        self.h = h
        self.w = w
        self.fovy = 2 * math.atan(2088.0 / (3443.915946 * 2))
        self.position = np.array([0.0, 0.0, 3.0]).astype(np.float32)
        self.target = np.array([0.0, 0.0, 0.0]).astype(np.float32)
        self.up = np.array([0.0, -1.0, 0.0]).astype(np.float32)
        # "arcball" state of the fork (glm.vec3 in the reference)
        self.camera_up = np.array([0.0, -1.0, 0.0], dtype=np.float32)
        self.camera_front = np.array([0.0, 0.0, -1.0], dtype=np.float32)
        self.camera_position = np.array([-3.0, 0.0, 1.5], dtype=np.float32)
        self.is_pose_dirty = True
        self.is_intrin_dirty = True
        self.target_dist = 3.0

    def get_view_matrix(self, arcball=True, front=None, pos=None, up=None, view=None):
        # /root/reference/util.py:58-70
        if arcball:
            if front is not None:
                if view is not None:
                    return np.array(view)
                pos = np.asarray(pos, dtype=np.float32)[:3]
                target = pos + np.asarray(front, dtype=np.float32)[:3]
                return look_at(pos, target, up)
            target = self.camera_position + self.camera_front
            return look_at(self.camera_position, target, self.camera_up)
        return look_at(self.position, self.target, self.up)

    def perspective(self, fov, aspect, near, far):
        # /root/reference/util.py:72-80
        f = 1 / np.tan(fov / 2.0)
        mat = np.zeros((4, 4))
        mat[0, 0] = f / aspect
        mat[1, 1] = f
        mat[2, 2] = -(far + near) / (far - near)
        mat[2, 3] = -(2.0 * far * near) / (far - near)
        mat[3, 2] = -1.0
        return mat

    def get_project_matrix(self):
        # /root/reference/util.py:82-105: near/far are hard-coded to 0.1/100 there
        aspect = self.w / self.h
        near = 0.1
        far = 100.0
        return np.array(self.perspective(self.fovy, aspect, near, far)).astype(np.float32)

    def get_htanfovxy_focal(self):
        # /root/reference/util.py:107-113
        htany = np.tan(self.fovy / 2)
        aspect_ratio = self.w / self.h
        htanx = htany * aspect_ratio
        focal = self.h / (2 * htany)
        return [htanx, htany, focal]

    def get_focal(self):
        return self.h / (2 * np.tan(self.fovy / 2))

    def flip_ground(self):
        self.up = -self.up
        self.is_pose_dirty = True

    def update_target_distance(self):
        _dir = self.target - self.position
        _dir = _dir / np.linalg.norm(_dir)
        self.target = self.position + _dir * self.target_dist

    def update_resolution(self, height, width):
        self.h = height
        self.w = width
        self.is_intrin_dirty = True


def pose_from_lookat(position, front, up) -> dict:
    """A pose dict with the keys the viewer's frame loop passes to the renderer
    (/root/reference/main.py:394-406): camera_front, camera_up, camera_position, camera_view."""
    position = np.asarray(position, dtype=np.float32)
    front = np.asarray(front, dtype=np.float32)
    up = np.asarray(up, dtype=np.float32)
    return {
        "camera_front": front,
        "camera_up": up,
        "camera_position": position,
        "camera_view": look_at(position, position + front, up),
    }
