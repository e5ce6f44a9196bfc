"""Ad-hoc (development build with -DGSR_DUP_TRACE only): globaltimer stamps of k_duplicate, first chunk of every CTA.
0 entry, 1 after griddepcontrol.wait, 2 ticket drawn, 3 loads + scan done, 4 chain prefix known, 5 emitted, 6 CTA end."""
import sys, ctypes as C
sys.path.insert(0, ".")
import numpy as np, torch
from gaussiansplattingviewer_b200 import _gsr, scenes
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
g = scenes.synthetic_gaussians(n, seed=0)
cam, pose = scenes.default_view(W, H)
r = B200Renderer(W, H)
r.update_gaussian_data(g); r.set_render_mod(3)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
for _ in range(6): r.draw()
torch.cuda.synchronize()
lib = _gsr.load()
buf = np.zeros(2048 * 8, dtype=np.uint64)
assert lib.gsr_debug_dup_trace(C.c_void_p(buf.ctypes.data), C.c_longlong(buf.size)) == 0
a = buf.reshape(2048, 8).astype(np.int64)
a = a[a[:, 6] > 0]
t0 = a[:, 1].min()
ph = (a[:, :7] - t0) / 1000.0
print(f"CTAs {len(a)}, chunks per CTA: mean {a[:, 7].mean():.2f} max {a[:, 7].max()} (CTAs without a chunk: {(a[:, 7] == 0).sum()}); kernel span after the dependency resolved: {ph[:, 6].max():.1f} us")
w = a[:, 7] > 0
d = np.diff(ph[w][:, 1:6], axis=1)
for i, l in enumerate(["ticket", "loads + scan", "chain look-back", "emit"]):
    print(f"   {l:16s} median {np.median(d[:, i]):6.2f} p90 {np.percentile(d[:, i], 90):6.2f} max {d[:, i].max():6.2f} us | ends at median {np.median(ph[w][:, i + 2]):6.2f} max {ph[w][:, i + 2].max():6.2f}")
print(f"   CTA end: median {np.median(ph[:, 6]):.2f} p90 {np.percentile(ph[:, 6], 90):.2f} max {ph[:, 6].max():.2f}; pdl wait of the CTAs: median {np.median(ph[:, 1] - ph[:, 0]):.2f}")
