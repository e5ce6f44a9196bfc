"""Writes build/lab/frame_*.bin for scripts/binlab.cu: tile rectangles and depth keys of the bench frames (from the CPU oracle)."""
import sys
sys.path.insert(0, "/root/repo")
import numpy as np
from gaussiansplattingviewer_b200 import scenes
from oracle import oracle as O
for tag, n, W, H in [("1080", 3_000_000, 1920, 1080), ("4k", 6_000_000, 3840, 2160)]:
    g = scenes.synthetic_gaussians(n, seed=0)
    cam, pose = scenes.default_view(W, H)
    ocam = O.make_camera(pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal(), pose["camera_position"], 1.0, 3, g.sh_dim)
    vs, st = O.preprocess(g, ocam, W, H)
    st = st.view(np.int32).reshape(-1, 8)
    ok = st[:, 3] == 1
    depth = st[:, 2].view(np.float32)
    tx0, tx1, ty0, ty1 = st[:, 4] >> 5, st[:, 5] >> 5, st[:, 6] >> 5, st[:, 7] >> 5
    rx = np.where(ok, tx0 | ((tx1 - tx0 + 1) << 16), 0).astype(np.uint32)
    ry = np.where(ok, ty0 | ((ty1 - ty0 + 1) << 16), 0).astype(np.uint32)
    dk = np.where(ok, depth.view(np.uint32), np.uint32(0xFFFFFFFF)).astype(np.uint32)
    TX, TY = (W + 31) // 32, (H + 31) // 32
    with open(f"/root/repo/build/lab/frame_{tag}.bin", "wb") as f:
        np.array([n, TX, TY, 0], dtype=np.int32).tofile(f)
        rx.tofile(f); ry.tofile(f); dk.tofile(f)
    print(tag, n, TX, TY, ok.sum())
