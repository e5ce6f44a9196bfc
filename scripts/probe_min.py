"""Ad-hoc: a few frames of a BASELINE config, for `ncu --metrics gpu__time_duration.sum` launch lists."""
import sys
sys.path.insert(0, ".")
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 3
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
r = B200Renderer(W, H)
import os
if os.environ.get("GSR_PROBE_NO_GRAPH"):  # kernel-by-kernel launches (ncu captures)
    from gaussiansplattingviewer_b200 import _gsr
    r.tune(_gsr.GSR_TUNE_GRAPH, 0)
r.update_gaussian_data(g)
r.set_render_mod(3)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
for _ in range(frames): r.draw()
import torch; torch.cuda.synchronize()
print("ok", r.stats().n_instances)
