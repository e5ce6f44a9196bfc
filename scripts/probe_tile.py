"""Ad-hoc: 16x16 vs 32x32 tiles — image equality and stage times (run on the GPU box)."""
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
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
imgs = {}
for mod in (3, -1, -3):
    r.set_render_mod(mod)
    for tile in (16, 32):
        r.tune(_gsr.GSR_TUNE_TILE, tile)
        imgs[(mod, tile)] = r.read_pixels_rgba32f().copy()
    a, b = imgs[(mod, 16)], imgs[(mod, 32)]
    print("mode", mod, "identical", np.array_equal(a.view(np.uint32), b.view(np.uint32)), "max abs diff", float(np.abs(a - b).max()), flush=True)
r.set_render_mod(3)
for tile in (16, 32, 16, 32):
    r.tune(_gsr.GSR_TUNE_TILE, tile)
    for _ in range(3): r.draw()
    acc = np.zeros(7); K = 20
    for _ in range(K):
        _, s = r.draw_timed()
        acc += [s.ms_preprocess, s.ms_sort_depth, s.ms_ranges, s.ms_duplicate, s.ms_sort_tiles, s.ms_blend, s.ms_total]
    acc /= K
    print("tile %d D %d | pre %.3f sortN %.3f scan %.3f dup %.3f sortD %.3f blend %.3f total %.3f" % ((tile, s.n_instances) + tuple(acc)), flush=True)
