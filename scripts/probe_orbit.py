"""Ad-hoc: per-stage times along the bench orbit (run on the GPU box)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
from gaussiansplattingviewer_b200.util import Camera
import bench
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
g = scenes.synthetic_gaussians(n, seed=0)
cam = Camera(H, W)
r = B200Renderer(W, H)
r.update_gaussian_data(g); r.set_render_mod(3); r.update_camera_intrin(cam)
acc = np.zeros(7); cnt = 0
for k in range(0, 55, 3):
    r.update_camera_pose(cam, True, bench.orbit_pose(k, 55))
    r.draw()
    _, s = r.draw_timed()
    v = [s.ms_preprocess, s.ms_sort_depth, s.ms_ranges, s.ms_duplicate, s.ms_sort_tiles, s.ms_blend, s.ms_total]
    acc += v; cnt += 1
    print(k, "V %d D %d pre %.3f sortN %.3f scan %.3f dup %.3f sortD %.3f blend %.3f tot %.3f" % ((s.n_visible, s.n_instances) + tuple(v)))
print("mean pre %.3f sortN %.3f scan %.3f dup %.3f sortD %.3f blend %.3f tot %.3f" % tuple(acc / cnt))
