"""Ad-hoc (development build with -DGSR_RS_TRACE only): where the radix passes of a multi-wave sort spend their time -
per-phase sums over ALL tiles of every CTA.  Pass 0 of gsr_sort_pairs lands in slot 3, the last pass in slot 4."""
import sys, ctypes as C
sys.path.insert(0, ".")
import numpy as np, torch
from gaussiansplattingviewer_b200 import _gsr
from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
m = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1 << 24
r = B200Renderer(64, 64)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(3):
    tk = torch.randint(-2**31, 2**31 - 1, (m,), dtype=torch.int32, device="cuda"); tv = torch.arange(m, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize(); e0.record(); r.sort_pairs(tk, tv); e1.record(); torch.cuda.synchronize()
print("pairs", m, "ms", round(e0.elapsed_time(e1), 4))
lib = _gsr.load()
buf = np.zeros(5 * 1024 * 8, dtype=np.uint64)
assert lib.gsr_debug_rs_acc(C.c_void_p(buf.ctypes.data), C.c_longlong(buf.size)) == 0
a = buf.reshape(5, 1024, 8).astype(np.int64)
tb = np.zeros(5 * 1024 * 10, dtype=np.uint64)
lib.gsr_debug_rs_trace(C.c_void_p(tb.ctypes.data), C.c_longlong(tb.size))
tb = tb.reshape(5, 1024, 10).astype(np.int64)
for s, name in ((3, "first pass"), (4, "last pass")):
    x = a[s]; x = x[x[:, 6] > 0]
    if len(x) == 0: continue
    tiles = x[:, 6]
    t0 = tb[s][:, 1][tb[s][:, 1] > 0].min()
    print(f"{name}: CTAs {len(x)}, tiles per CTA {tiles.mean():.2f}; kernel span {(x[:, 7].max() - t0) / 1000:.1f} us")
    lab = ["ticket", "clear status", "sub-tile 0", "sub-tile 1", "prefix wait", "scatter"]
    per = x[:, :6] / tiles[:, None] / 1000.0
    tot = per.sum(axis=1)
    for i, l in enumerate(lab):
        print(f"   {l:14s} per tile: mean {per[:, i].mean():6.2f} us ({100 * per[:, i].mean() / tot.mean():4.1f} %)")
    print(f"   per tile total {tot.mean():.2f} us; x tiles per CTA = {(tot * tiles).mean():.1f} us")
