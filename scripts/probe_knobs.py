"""Ad-hoc: stage times under gsr_tune knobs given as key=value pairs on the command line (run on the GPU box)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from gaussiansplattingviewer_b200 import scenes, _gsr
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
n = 3_000_000
W, H = 1920, 1080
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
r = B200Renderer(W, H)
r.update_gaussian_data(g)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
configs = [c for c in sys.argv[1:]] or ["-"]
for cfg in configs + configs:
    pairs = [] if cfg == "-" else [tuple(int(x) for x in kv.split("=")) for kv in cfg.split(",")]
    for k, v in pairs: r.tune(k, v)
    for _ in range(3): r.draw()
    acc = np.zeros(7); K = 20
    for _ in range(K):
        _, s = r.draw_timed()
        acc += [s.ms_preprocess, s.ms_sort_depth, s.ms_ranges, s.ms_duplicate, s.ms_sort_tiles, s.ms_blend, s.ms_total]
    acc /= K
    print("%-12s | pre %.3f sortN %.3f scan %.3f dup %.3f sortD %.3f blend %.3f total %.3f" % ((cfg,) + tuple(acc)), flush=True)
