"""Ad-hoc: per-stage CUDA-event times (gsr_render_timed) and the standalone pair sort, for A/B runs of kernel variants."""
import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
r = B200Renderer(W, H)
r.update_gaussian_data(g); r.set_render_mod(3)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
for _ in range(5): r.draw()
acc = np.zeros(7); K = 20
for _ in range(K):
    _, s = r.draw_timed()
    acc += [s.ms_preprocess, s.ms_sort_depth, s.ms_ranges, s.ms_duplicate, s.ms_sort_tiles, s.ms_blend, s.ms_total]
print("stage ms", dict(zip("pre depth scan dup tile blend total".split(), np.round(acc / K, 4))))
torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): r.draw()
e1.record(); torch.cuda.synchronize()
print("frame ms", round(e0.elapsed_time(e1) / 50, 4))
m = 1 << 24
best = 1e9
for _ in range(5):
    tk = torch.randint(-2**31, 2**31 - 1, (m,), dtype=torch.int32, device="cuda"); tv = torch.arange(m, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize(); e0.record(); r.sort_pairs(tk, tv); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print("pairs 2^24 ms", round(best, 4), "Gkeys/s", round(m / best / 1e6, 2))
