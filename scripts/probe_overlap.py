"""Ad-hoc: frames/s of one handle vs two handles rendering alternate poses on two streams (run on the GPU box)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
views, campos = scenes.orbit_poses(64)
rs = []
for k in range(3):
    r = B200Renderer(W, H)
    r.update_gaussian_data(g)
    r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
    rs.append(r)
streams = [torch.cuda.Stream() for _ in range(3)]
def run(nh, K=128):
    for w in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(K):
            r = rs[i % nh]
            r._view = views[i % 64]; r._campos = campos[i % 64]
            with torch.cuda.stream(streams[i % nh]):
                r.draw()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    return K / dt
for nh in (1, 2, 3, 1, 2, 3):
    print("handles", nh, "fps %.1f" % run(nh), flush=True)
