"""Ad-hoc: sync vs async RGB8 read-back vs the float frame, per pose (run on the GPU box)."""
import sys, numpy as np
sys.path.insert(0, ".")
from gaussiansplattingviewer_b200 import scenes, _gsr, util
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
g = scenes.synthetic_gaussians(40000, seed=27)
W, H = 480, 270
cam0, pose0 = scenes.default_view(W, H)
r = B200Renderer(W, H)
r.update_gaussian_data(g); r.set_render_mod(3)
r.update_camera_intrin(cam0); r.update_camera_pose(cam0, True, pose0)
views, pos = scenes.orbit_poses(7)
cam = util.Camera(H, W)
def pose(i):
    return {"camera_front": np.zeros(3, np.float32), "camera_up": np.zeros(3, np.float32), "camera_position": pos[i], "camera_view": views[i]}
for i in range(3):
    r.update_camera_pose(cam, True, pose(i))
    f = r.read_pixels_rgba32f()[..., :3]
    ref = np.rint(np.clip(f, 0, 1) * 255).astype(np.uint8)
    s = r.read_pixels_rgb8().copy()
    a = r.read_pixels_rgb8_async(); r.sync_host(); a = a.copy()
    bad_s = (s != ref).any(axis=2); bad_a = (a != ref).any(axis=2)
    print(i, "sync bad px", int(bad_s.sum()), "rows", np.flatnonzero(bad_s.any(axis=1))[:6], "| async bad px", int(bad_a.sum()),
          "rows", np.flatnonzero(bad_a.any(axis=1))[:6], "cols", np.flatnonzero(bad_a.any(axis=0))[:6], flush=True)
