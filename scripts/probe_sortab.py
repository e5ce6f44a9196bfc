"""Ad-hoc A/B of library variants on the sorts: pairs 2^24 / 2^22, and the 6M / 4K frame's tile sort."""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
from gaussiansplattingviewer_b200 import scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
tag = sys.argv[1] if len(sys.argv) > 1 else ""
r = B200Renderer(64, 64)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
out = []
for m in (1 << 24, 1 << 22):
    best = 1e9
    for _ in range(6):
        tk = torch.randint(-2**31, 2**31 - 1, (m,), dtype=torch.int32, device="cuda"); tv = torch.arange(m, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize(); e0.record(); r.sort_pairs(tk, tv); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    out.append(f"pairs 2^{m.bit_length() - 1} {best:.4f} ms ({m / best / 1e6:.2f} Gkeys/s)")
del r
g = scenes.synthetic_gaussians(6_000_000, seed=0)
cam, pose = scenes.default_view(3840, 2160)
r = B200Renderer(3840, 2160)
r.update_gaussian_data(g); r.set_render_mod(3)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
for _ in range(5): r.draw()
acc = np.zeros(2)
for _ in range(20):
    _, s = r.draw_timed(); acc += [s.ms_sort_depth, s.ms_sort_tiles]
out.append(f"6M/4K depth {acc[0] / 20:.4f} tile {acc[1] / 20:.4f} ms")
print(tag, "; ".join(out), flush=True)
