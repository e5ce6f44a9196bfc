"""COLMAP ``images.txt`` / ``cameras.txt`` -> stereo pose batches for the capture loop.

Restates the pose part of the viewer's main loop (/root/reference/main.py): the text parsing of
``main(...)`` (:602-632), ``quaternion_to_rotation_matrix`` (:169-182), ``create_look_at_from_colmap``
(:197-217) and ``load_camera_positions`` (:275-407), which turns one ``images.txt`` row into the LEFT pose
and a RIGHT pose whose view matrix is the left one pre-multiplied by an x translation of ``baseline`` =
-0.5 in view space (:376-380), i.e. a camera 0.5 units to the right on the same image plane.

The reference builds the view with ``glm.lookAtRH`` on ``glm.vec3`` (float32) inputs; PyGLM is not
installed here, so ``util.look_at`` (glm's formula in float32) stands in.  Parity: the NumPy functions are
pinned bit for bit to the reference's own code (tests/golden/make_colmap_golden.py cuts them out of main.py
and runs them); ``load_camera_positions`` is pinned to the reference's call sequence run on a NumPy shim of
GLM's float32 formulas (1 ulp), not to a PyGLM binary.

Everything here is host-side NumPy: a few hundred 4x4 matrices per scene.  The arrays it returns feed
``B200Renderer.capture_stereo`` / ``draw_batch`` (C ABI: gsr_capture_stereo_async / gsr_render_batch).
"""
from __future__ import annotations

import os

import numpy as np

from . import util

BASELINE = -0.5  # main.py:280


def read_images_txt(path):
    """Rows of ``images.txt`` as lists of strings ``[IMAGE_ID, QW, QX, QY, QZ, TX, TY, TZ, CAMERA_ID, NAME]``.
    Like main.py:608-622: comment lines are skipped, every second remaining line (the 2D points) is dropped,
    and a pose line must split into exactly 10 fields (the reference's tuple unpacking raises otherwise)."""
    poses = []
    with open(path, "r") as f:
        line_no = 0
        for line in f.readlines():
            if line.startswith("#"):
                continue
            if line_no % 2 == 1:
                line_no += 1
                continue
            elements = line.split()
            if len(elements) != 10:
                raise ValueError(f"{path}: expected 10 fields in a pose line, got {len(elements)}")
            poses.append(elements)
            line_no += 1
    return poses


def read_cameras_txt(path):
    """Last camera of ``cameras.txt`` as a dict (id, model, width, height, fx, fy, cx, cy), main.py:625-632.
    The reference parses these and then overrides the resolution with 1160x522 (:634-636)."""
    cam = None
    with open(path, "r") as f:
        for line in f.readlines():
            if line.startswith("#"):
                continue
            e = line.split()
            if len(e) < 8:
                continue
            cam = {"id": int(e[0]), "model": e[1], "width": int(e[2]), "height": int(e[3]),
                   "fx": float(e[4]), "fy": float(e[5]), "cx": float(e[6]), "cy": float(e[7])}
    return cam


def quaternion_to_rotation_matrix(qw, qx, qy, qz):
    # main.py:169-182 (float64)
    q = np.array([qw, qx, qy, qz], dtype=np.float64)
    norm = np.linalg.norm(q)
    if norm == 0:
        raise ValueError("Cannot normalize a zero-norm quaternion.")
    qw, qx, qy, qz = q / norm
    return np.array([
        [1 - 2 * qy ** 2 - 2 * qz ** 2, 2 * qx * qy - 2 * qz * qw, 2 * qx * qz + 2 * qy * qw],
        [2 * qx * qy + 2 * qz * qw, 1 - 2 * qx ** 2 - 2 * qz ** 2, 2 * qy * qz - 2 * qx * qw],
        [2 * qx * qz - 2 * qy * qw, 2 * qy * qz + 2 * qx * qw, 1 - 2 * qx ** 2 - 2 * qy ** 2]])


def create_look_at_from_colmap(tx, ty, tz, qw, qx, qy, qz):
    """(camera_pos, center_point, world_up), main.py:197-217: the camera sits at -T, its axes are the columns
    of R^T with z flipped, it looks along the rotated (0,0,-1) with the rotated (0,-1,0) as up."""
    camera_pos = np.array([-tx, -ty, -tz], dtype=np.float64)
    rot = quaternion_to_rotation_matrix(qw, qx, qy, qz).T
    rot_flip = rot @ np.diag([1.0, 1.0, -1.0])
    world_forward = rot_flip @ np.array([0.0, 0.0, -1.0])
    world_up = rot_flip @ np.array([0.0, -1.0, 0.0])
    return camera_pos, camera_pos + world_forward, world_up


def right_view(view_left, baseline=BASELINE):
    """``T * viewMat`` with ``T[3,0] = baseline`` (glm column 3, row 0), main.py:376-380: in maths layout the
    x translation of the view matrix grows by ``baseline``; rotation and the y/z rows are untouched, so every
    point keeps its view-space depth."""
    v = np.array(view_left, dtype=np.float32, copy=True)
    v[0, 3] = np.float32(v[0, 3] + np.float32(baseline))
    return v


def camera_position_of(view):
    """``glm.inverse(view)[3]`` (main.py:403): the eye of a rigid view matrix, -R^T t."""
    v = np.asarray(view, dtype=np.float32)
    return (-(v[:3, :3].T @ v[:3, 3])).astype(np.float32)


def load_camera_positions(camera_pose, baseline=BASELINE):
    """One ``images.txt`` row -> (pose_left, pose_right) dicts with the keys the renderer's
    ``sort_and_update`` / ``update_camera_pose`` read (main.py:394-406).  ``camera_front`` / ``camera_up`` are
    the module-level vectors of the reference (:160-161, up replaced by the look-at up at :322)."""
    qw, qx, qy, qz = (float(camera_pose[i]) for i in (1, 2, 3, 4))
    x, y, z = (float(camera_pose[i]) for i in (5, 6, 7))
    camera_pos, center_point, up = create_look_at_from_colmap(x, y, z, qw, qx, qy, qz)
    view_left = util.look_at(camera_pos, center_point, up)  # glm.lookAtRH on float32 vectors (:324)
    view_r = right_view(view_left, baseline)
    front = np.array([0.0, 0.0, 1.0], dtype=np.float32)
    up32 = np.asarray(up, dtype=np.float32)
    pose_left = {"camera_front": front, "camera_up": up32, "camera_position": camera_pos.astype(np.float32),
                 "camera_view": view_left, "image_id": int(camera_pose[0])}
    pose_right = {"camera_front": front, "camera_up": up32, "camera_position": camera_position_of(view_r),
                  "camera_view": view_r, "image_id": int(camera_pose[0])}
    return pose_left, pose_right


def stereo_pose_batch(colmap_dir_or_rows, baseline=BASELINE):
    """All poses of a COLMAP model as arrays: views_left (n,4,4), campos_left (n,3), views_right (n,4,4),
    campos_right (n,3), image_ids (n,) — the batch a capture run walks (main.py:793-917, one scene of
    render_all.py:14-35)."""
    rows = colmap_dir_or_rows
    if isinstance(rows, (str, os.PathLike)):
        rows = read_images_txt(os.path.join(rows, "images.txt"))
    n = len(rows)
    vl = np.zeros((n, 4, 4), np.float32)
    vr = np.zeros((n, 4, 4), np.float32)
    cl = np.zeros((n, 3), np.float32)
    cr = np.zeros((n, 3), np.float32)
    ids = np.zeros(n, np.int64)
    for i, row in enumerate(rows):
        pl, pr = load_camera_positions(row, baseline)
        vl[i], cl[i] = pl["camera_view"], pl["camera_position"]
        vr[i], cr[i] = pr["camera_view"], pr["camera_position"]
        ids[i] = pl["image_id"]
    return vl, cl, vr, cr, ids
