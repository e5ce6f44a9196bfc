"""Ad-hoc A/B of library variants (build/var/libgsr_*.so copied over csrc/libgsr.so on the GPU box by the caller):
blend stage, frame time and an image hash at two workloads."""
import sys, zlib
sys.path.insert(0, ".")
import numpy as np, torch
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
tag = sys.argv[1] if len(sys.argv) > 1 else ""
for n, W, H in [(3_000_000, 1920, 1080), (6_000_000, 3840, 2160)]:
    g = scenes.synthetic_gaussians(n, seed=0)
    cam, pose = scenes.default_view(W, H)
    r = B200Renderer(W, H)
    r.update_gaussian_data(g); r.set_render_mod(3)
    r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
    for _ in range(5): img = r.draw()
    h = zlib.crc32(img.cpu().numpy().tobytes())
    K = 30; acc = np.zeros(5)
    for _ in range(K):
        _, s = r.draw_timed(); acc += [s.ms_preprocess, s.ms_sort_tiles, s.ms_blend, s.ms_sort_depth, s.ms_duplicate]
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(60): r.draw()
    e1.record(); torch.cuda.synchronize()
    a = acc / K
    print(f"{tag:8s} {n} {W}x{H}: pre {a[0]:.4f} depth {a[3]:.4f} dup {a[4]:.4f} tile {a[1]:.4f} blend {a[2]:.4f} frame {e0.elapsed_time(e1) / 60:.4f} ms  crc {h:08x}", flush=True)
    del r
