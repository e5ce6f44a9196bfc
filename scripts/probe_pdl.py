"""Ad-hoc: untimed frames/s with and without programmatic dependent launch (run on the GPU box)."""
import sys
sys.path.insert(0, ".")
import torch
from gaussiansplattingviewer_b200 import scenes, _gsr
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
n, W, H = 3_000_000, 1920, 1080
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
views, campos = scenes.orbit_poses(64)
r = B200Renderer(W, H)
r.update_gaussian_data(g)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
for v in [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else "0,1,0,1".split(","))]:
    r.tune(_gsr.GSR_TUNE_PDL, v)
    for i in range(8):
        r._view = views[i]; r._campos = campos[i]; r.draw()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 256
    e0.record()
    for i in range(K):
        r._view = views[i % 64]; r._campos = campos[i % 64]; r.draw()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    print("pdl=%d | %.4f ms/frame  %.1f frames/s" % (v, ms, 1000.0 / ms), flush=True)
