"""Ad-hoc: is the first synchronous RGB8 frame of a fresh process equal to later ones?  One knob setting per
subprocess (run on the GPU box)."""
import subprocess, sys
CHILD = r'''
import sys, numpy as np
sys.path.insert(0, ".")
from gaussiansplattingviewer_b200 import scenes, _gsr, util
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
knobs = [tuple(int(x) for x in kv.split("=")) for kv in sys.argv[1].split(",") if kv]
g = scenes.synthetic_gaussians(40000, seed=27)
W, H = 480, 270
cam, pose = scenes.default_view(W, H)
r = B200Renderer(W, H)
for k, v in knobs: r.tune(k, v)
r.update_gaussian_data(g); r.set_render_mod(3)
r.update_camera_intrin(cam); r.update_camera_pose(cam, True, pose)
a = r.read_pixels_rgb8().copy()
b = r.read_pixels_rgb8().copy()
c = r.read_pixels_rgb8().copy()
print(sys.argv[1] or "-", "first==second", np.array_equal(a, b), "second==third", np.array_equal(b, c),
      "diff px", int((a != b).any(axis=2).sum()), int((b != c).any(axis=2).sum()))
'''
for cfg in ["", "10=0", "9=0", "8=0", "6=0", "10=0,9=0,8=0,6=0"]:
    subprocess.run([sys.executable, "-c", CHILD, cfg])
