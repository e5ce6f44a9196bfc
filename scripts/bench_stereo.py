"""Stereo capture throughput (left RGB8 + disparity u16 + right RGB8 per pose, main.py:844-917).

    python scripts/bench_stereo.py [--n 3000000] [--w 1920 --h 1080] [--poses 60]

One JSON line: poses/s through B200Renderer.capture_stereo (pinned host buffers, copies overlapped) with the
shared depth order on and off, against three independent read_pixels_* calls per pose (what a straight port of
the reference loop does), and against the frozen order (auto sort off).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=3_000_000)
    ap.add_argument("--w", type=int, default=1920)
    ap.add_argument("--h", type=int, default=1080)
    ap.add_argument("--poses", type=int, default=60)
    a = ap.parse_args()
    import torch
    from gaussiansplattingviewer_b200 import colmap, scenes, util
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer

    g = scenes.synthetic_gaussians(a.n, seed=0)
    r = B200Renderer(a.w, a.h)
    r.update_gaussian_data(g)
    cam = util.Camera(a.h, a.w)
    r.update_camera_intrin(cam)
    views, campos = scenes.orbit_poses(a.poses)
    poses = []
    for v, c in zip(views, campos):
        front = np.array([0, 0, 1], np.float32)
        up = np.array([0, -1, 0], np.float32)
        pl = {"camera_view": v, "camera_position": c, "camera_front": front, "camera_up": up}
        vr = colmap.right_view(v)
        pr = {"camera_view": vr, "camera_position": colmap.camera_position_of(vr), "camera_front": front, "camera_up": up}
        poses.append((pl, pr))

    def run_capture(share):
        for pl, pr in poses[:3]:
            r.capture_stereo(pl, pr, share_order=share)
        r.sync_host()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for pl, pr in poses:
            r.capture_stereo(pl, pr, share_order=share)
        r.sync_host()
        torch.cuda.synchronize()
        return len(poses) / (time.perf_counter() - t0)

    def run_independent():
        def one(pl, pr):
            r.set_render_mod(3)
            r.update_camera_pose(cam, True, pl)
            r.read_pixels_rgb8()
            r.set_render_mod(-1)
            r.read_pixels_r32f()
            r.set_render_mod(3)
            r.update_camera_pose(cam, True, pr)
            r.read_pixels_rgb8()
        for pl, pr in poses[:3]:
            one(pl, pr)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for pl, pr in poses:
            one(pl, pr)
        torch.cuda.synchronize()
        r.set_render_mod(3)
        return len(poses) / (time.perf_counter() - t0)

    out = {"metric": "stereo poses/s (left RGB8 + disparity u16 + right RGB8)", "n_gaussians": a.n, "width": a.w,
           "height": a.h, "poses": len(poses)}
    out["capture_shared_order"] = run_capture(True)
    out["capture_three_sorts"] = run_capture(False)
    out["three_synchronous_readbacks"] = run_independent()
    r.set_auto_sort(False)
    r.sort_and_update(cam, True, poses[0][0])
    out["capture_frozen_order"] = run_capture(True)
    r.set_auto_sort(True)
    out["d2h_bytes_per_pose"] = a.w * a.h * 8
    print(json.dumps(out))


if __name__ == "__main__":
    main()
