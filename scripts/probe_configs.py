"""Per-stage times and untimed frames/s for the BASELINE.json configurations (run on the GPU box)."""
import sys
import numpy as np
sys.path.insert(0, ".")
import torch
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer

CASES = [(1_000_000, 1920, 1080, 3), (3_000_000, 1920, 1080, 3), (6_000_000, 3840, 2160, 3),
         (6_000_000, 3840, 2160, -1), (6_000_000, 3840, 2160, -3)]
last = None
for n, W, H, mod in CASES:
    if last is None or last[0] != (n, W, H):
        g = scenes.synthetic_gaussians(n, seed=0)
        cam, pose = scenes.default_view(W, H)
        r = B200Renderer(W, H)
        r.update_gaussian_data(g)
        r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
        last = ((n, W, H), r)
    r = last[1]
    r.set_render_mod(mod)
    for _ in range(3): r.draw()
    acc = np.zeros(7); K = 10
    for _ in range(K):
        _, s = r.draw_timed()
        acc += [s.ms_preprocess, s.ms_sort_depth, s.ms_ranges, s.ms_duplicate, s.ms_sort_tiles, s.ms_blend, s.ms_total]
    acc /= K
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): r.draw()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    st = r.stats()
    print("n=%d %dx%d mod=%d | pre %.3f sortN %.3f scan %.3f dup %.3f sortD %.3f blend %.3f | %.3f ms/frame %.0f frames/s | visible %d instances %d"
          % ((n, W, H, mod) + tuple(acc[:6]) + (ms, 1000.0 / ms, st.n_visible, st.n_instances)), flush=True)
