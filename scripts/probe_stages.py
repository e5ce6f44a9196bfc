"""Ad-hoc: per-stage times of the frame pipeline at a BASELINE config (run on the GPU box)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
mod = int(sys.argv[4]) if len(sys.argv) > 4 else 3
keys64 = len(sys.argv) > 5 and sys.argv[5] == "k64"
t = time.time(); g = scenes.synthetic_gaussians(n, seed=0); print("gen", round(time.time() - t, 2), "s")
cam, pose = scenes.default_view(W, H)
r = B200Renderer(W, H, keys64=keys64)
t = time.time(); r.update_gaussian_data(g); print("upload", round(time.time() - t, 2), "s")
r.set_render_mod(mod)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
for _ in range(3): r.draw()
acc = np.zeros(7)
K = 10
for _ in range(K):
    _, s = r.draw_timed()
    acc += [s.ms_preprocess, s.ms_sort_depth, s.ms_duplicate, s.ms_sort_tiles, s.ms_ranges, s.ms_blend, s.ms_total]
acc /= K
print("N", n, "V", s.n_visible, "D", s.n_instances, "cap", s.capacity, "ovf", s.overflow)
print("ms: pre %.3f sortN %.3f dup %.3f sortD %.3f ranges %.3f blend %.3f total %.3f" % tuple(acc))
import torch
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(20): r.draw()
e1.record(); torch.cuda.synchronize()
print("untimed loop ms/frame %.3f  fps %.1f" % (e0.elapsed_time(e1) / 20, 20e3 / e0.elapsed_time(e1)))
print("pairs visited/contributed", r.count_pairs(), "peaks", r.measure_peaks())
