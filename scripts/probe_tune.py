"""Ad-hoc: per-stage times for each kernel-variant knob (run on the GPU box)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from gaussiansplattingviewer_b200 import scenes, _gsr
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
r = B200Renderer(W, H)
r.update_gaussian_data(g)
r.set_render_mod(3)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)


def run(tag):
    for _ in range(3): r.draw()
    acc = np.zeros(7); K = 10
    for _ in range(K):
        _, s = r.draw_timed()
        acc += [s.ms_preprocess, s.ms_sort_depth, s.ms_ranges, s.ms_duplicate, s.ms_sort_tiles, s.ms_blend, s.ms_total]
    acc /= K
    print("%-22s pre %.3f sortN %.3f scan %.3f dup %.3f sortD %.3f blend %.3f total %.3f" % ((tag,) + tuple(acc)), flush=True)


run("default")
T = _gsr
knob = sys.argv[4] if len(sys.argv) > 4 else "GSR_TUNE_SORT_PREFETCH"
vals = [int(v) for v in sys.argv[5].split(",")] if len(sys.argv) > 5 else [0, 1, 0, 1]
for v in vals:
    r.tune(getattr(T, knob), v); run(f"{knob[9:].lower()}={v}")
