"""Ad-hoc: untimed frames/s over an orbit under gsr_tune settings given as key=value[,key=value] (GPU box)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from gaussiansplattingviewer_b200 import scenes, _gsr
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
n = 3_000_000
W, H = 1920, 1080
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
views, campos = scenes.orbit_poses(64)
r = B200Renderer(W, H)
r.update_gaussian_data(g)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
configs = sys.argv[1:] or ["-"]
for cfg in configs + configs:
    pairs = [] if cfg == "-" else [tuple(int(x) for x in kv.split("=")) for kv in cfg.split(",")]
    for k, v in pairs: r.tune(k, v)
    for i in range(8):
        r._view = views[i]; r._campos = campos[i]; r.draw()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 128
    e0.record()
    for i in range(K):
        r._view = views[i % 64]; r._campos = campos[i % 64]; r.draw()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    print("%-14s | %.4f ms/frame  %.1f frames/s" % (cfg, ms, 1000.0 / ms), flush=True)
