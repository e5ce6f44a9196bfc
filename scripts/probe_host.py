"""Ad-hoc: host time to enqueue a frame (Python -> ctypes -> driver), and frame time, with and without the captured graph."""
import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from gaussiansplattingviewer_b200 import scenes, _gsr
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
for n, W, H in [(3_000_000, 1920, 1080), (4, 1280, 720)]:
    g = scenes.synthetic_gaussians(n, seed=0)
    cam, pose = scenes.default_view(W, H)
    for graph in (1, 0):
        r = B200Renderer(W, H)
        r.tune(_gsr.GSR_TUNE_GRAPH, graph)
        r.update_gaussian_data(g); r.set_render_mod(3)
        r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
        for _ in range(10): r.draw()
        torch.cuda.synchronize()
        K = 300
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record()
        for _ in range(K): r.draw()
        t1 = time.perf_counter(); e1.record(); torch.cuda.synchronize()
        print(f"n={n} {W}x{H} graph={graph}: host enqueue {1e6 * (t1 - t0) / K:.1f} us/frame, frame {e0.elapsed_time(e1) / K:.4f} ms")
        del r
