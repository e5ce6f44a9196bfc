"""Ad-hoc (development build with -DGSR_RS_TRACE only): per-phase globaltimer stamps of the radix passes of a frame.
Phases per CTA (first super-tile): 0 entry, 1 after griddepcontrol.wait, 2 ticket, 3 status rows cleared, 4 sub-tile 0
parked, 5 sub-tile 1 parked, 6 prefix known, 7 scattered."""
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
buf = np.zeros(5 * 1024 * 10, dtype=np.uint64)
rc = lib.gsr_debug_rs_trace(C.c_void_p(buf.ctypes.data), C.c_longlong(buf.size))
assert rc == 0, rc
tr = buf.reshape(5, 1024, 10).astype(np.int64)
buf2 = np.zeros(5 * 1024 * 10, dtype=np.uint64)
tr2 = None
if hasattr(lib, "gsr_debug_rs_trace2") and lib.gsr_debug_rs_trace2(C.c_void_p(buf2.ctypes.data), C.c_longlong(buf2.size)) == 0:
    tr2 = buf2.reshape(5, 1024, 10).astype(np.int64)
names = ["depth p0 (drop)", "depth p1", "depth p2", "tile p0", "tile p1"]
for s in range(5):
    a = tr[s]
    used = a[:, 7] > 0
    a = a[used]
    if len(a) == 0: continue
    t0 = a[:, 0].min()
    ph = (a[:, :8] - t0) / 1000.0
    d = np.diff(ph, axis=1)
    print(f"{names[s]}: CTAs with a tile {len(a)}, tiles {int(a[0, 9])}; kernel span {ph[:, 7].max():.1f} us")
    lab = ["pdl wait", "hist scan+ticket", "clear status", "sub-tile 0", "sub-tile 1", "prefix wait", "scatter"]
    for i, l in enumerate(lab):
        print(f"   {l:18s} median {np.median(d[:, i]):6.2f}  p90 {np.percentile(d[:, i], 90):6.2f}  max {d[:, i].max():6.2f} us"
              f"   | phase ends at: median {np.median(ph[:, i + 1]):6.2f} max {ph[:, i + 1].max():6.2f}")
    tiles = a[:, 8] & 0xffffffff
    o = np.argsort(tiles)
    # when does tile k know its prefix, relative to when the LAST tile in front of it finished counting?
    end5 = ph[o, 5]; end6 = ph[o, 6]
    run = np.maximum.accumulate(end5)
    lag = end6[1:] - run[:-1]
    print(f"   prefix known after the slowest predecessor counted: median {np.median(lag):.2f} p90 {np.percentile(lag, 90):.2f} max {lag.max():.2f} us;"
          f" counted (phase 5) spread: min {end5.min():.2f} median {np.median(end5):.2f} max {end5.max():.2f}")
    if tr2 is not None:
        b = tr2[s][used]
        q = (b[:, :5] - t0) / 1000.0
        cnt_all = ph[:, 5].max()
        lastg = b[:, 7] == 1
        print(f"   inside tile_prefix (us after the LAST tile of the pass had counted = {cnt_all:.2f}): own-group rows done median {np.median(q[:, 1]) - cnt_all:6.2f} max {q[:, 1].max() - cnt_all:6.2f};"
              f" (group publishers: median {np.median(q[lastg, 1]) - cnt_all:6.2f} max {q[lastg, 1].max() - cnt_all:6.2f});"
              f" first group step back median {np.median(q[:, 2][b[:, 2] > 0]) - cnt_all:6.2f}; exit median {np.median(q[:, 4]) - cnt_all:6.2f} max {q[:, 4].max() - cnt_all:6.2f}")
        d_own = q[:, 1] - q[:, 0]; d_grp = q[:, 4] - q[:, 1]
        rr = (b[:, 8] if False else (a[:, 8] & 0xffffffff)) % 16
        print(f"   own-group phase: median {np.median(d_own):.2f} p90 {np.percentile(d_own, 90):.2f} max {d_own.max():.2f} us (r=0 tiles: {np.median(d_own[rr == 0]):.2f}, r=15: {np.median(d_own[rr == 15]):.2f});"
              f" group phase: median {np.median(d_grp):.2f} p90 {np.percentile(d_grp, 90):.2f} max {d_grp.max():.2f}; entry after counted: {np.median(q[:, 0] - ph[:, 5]):.2f}")
        if s == 1:
            tl = a[:, 8] & 0xffffffff
            for k in list(range(0, 48)):
                i = np.where(tl == k)[0]
                if len(i):
                    i = i[0]
                    print(f"      tile {k:3d}: counted {ph[i, 5]:6.2f} entry {q[i, 0]:6.2f} own done {q[i, 1]:6.2f} (steps {b[i, 5] & 0xffffffff} waits {b[i, 5] >> 32}) first grp step {q[i, 2] if b[i, 2] > 0 else 0:6.2f} exit {q[i, 4]:6.2f} (steps {b[i, 6] & 0xffffffff} waits {b[i, 6] >> 32})")
        print(f"   steps/waits own group: {np.mean(b[:, 5] & 0xffffffff):.2f} / {np.mean(b[:, 5] >> 32):.2f}; groups: {np.mean(b[:, 6] & 0xffffffff):.2f} / {np.mean(b[:, 6] >> 32):.2f}")
    if s in (1, 3):
        for k in range(0, len(o), max(1, len(o) // 28)):
            i = o[k]
            print(f"      tile {int(tiles[i]):4d} sm {int(a[i, 8] >> 32):3d}: start {ph[i, 1]:6.2f} counted {ph[i, 5]:6.2f} prefix {ph[i, 6]:6.2f} done {ph[i, 7]:6.2f}")
