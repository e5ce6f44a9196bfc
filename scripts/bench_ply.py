"""PLY decode: device kernel vs the host load_ply, on a synthetic 3DGS point cloud (62 float props).

    python scripts/bench_ply.py [--n 3000000] [--cpu-n 300000] [--reps 5]

Prints one JSON line: device-resident decode GB/s (algorithmic bytes = 248 read + 236 written per vertex)
against the measured HBM peak, the end-to-end time from pinned host rows (upload overlapped with the
decode), the serial float32 running-sum kernel for the mean, and the host load_ply on a bounded sample.
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=3_000_000)
    ap.add_argument("--cpu-n", type=int, default=300_000)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    import torch
    from gaussiansplattingviewer_b200 import util_gau
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer

    n = a.n
    rng = np.random.default_rng(0)
    rows = np.empty((n, 62), dtype=np.float32)
    rows[:] = rng.standard_normal((n, 62), dtype=np.float32)
    names = ["x", "y", "z", "nx", "ny", "nz"] + util_gau.PLY_COLUMNS[3:]
    props = [(nm, "f4") for nm in names]
    row_bytes, offs = util_gau.ply_row_layout(props)
    r = B200Renderer(64, 64)
    pinned = torch.empty(n * row_bytes, dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[:] = rows.view(np.uint8).ravel()
    dev = pinned.cuda()
    torch.cuda.synchronize()

    def timed(fn):
        ts = []
        for _ in range(a.reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        return min(ts), float(np.median(ts))

    t_dev = timed(lambda: r.load_ply_rows(dev, n, row_bytes, offs))
    t_host = timed(lambda: r.load_ply_rows(pinned.numpy(), n, row_bytes, offs))

    # host load_ply on a bounded sample of the same rows
    m = min(a.cpu_n, n)
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "s.ply")
        hdr = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % m
        hdr += "".join(f"property float {nm}\n" for nm in names) + "end_header\n"
        with open(p, "wb") as f:
            f.write(hdr.encode())
            f.write(rows[:m].tobytes())
        t0 = time.perf_counter()
        with np.errstate(all="ignore"):
            util_gau.load_ply(p)
        t_cpu = time.perf_counter() - t0
        t0 = time.perf_counter()
        r.load_ply(p)
        t_file = time.perf_counter() - t0
    # the whole file, end to end: B200Renderer.load_ply(path) (header parse + mapped body + chunked upload + decode)
    # with the page cache warm and, where the box lets us drop it, cold; the NumPy port beside it
    full = {}
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "full.ply")
        hdr = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n
        hdr += "".join(f"property float {nm}\n" for nm in names) + "end_header\n"
        with open(p, "wb") as f:
            f.write(hdr.encode())
            f.write(rows.tobytes())
        r.load_ply(p)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); r.load_ply(p); ts.append(time.perf_counter() - t0)
        full["device_from_file_warm_s"] = min(ts)
        try:
            os.sync()
            with open("/proc/sys/vm/drop_caches", "w") as f:
                f.write("3\n")
            t0 = time.perf_counter(); r.load_ply(p); full["device_from_file_cold_s"] = time.perf_counter() - t0
        except Exception as e:
            full["device_from_file_cold_s"] = None
            full["cold_note"] = f"page cache could not be dropped here ({type(e).__name__})"
        t0 = time.perf_counter()
        with np.errstate(all="ignore"):
            util_gau.load_ply(p)
        full["cpu_load_ply_s"] = time.perf_counter() - t0
        full["n"] = n
        full["file_bytes"] = os.path.getsize(p)
    peaks = {}
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    alg = n * (row_bytes + 236)
    line = {
        "metric": "PLY decode (load_ply body) of N 3DGS vertices", "n": n, "row_bytes": row_bytes,
        "device_resident": {"ms_min": t_dev[0] * 1e3, "ms_median": t_dev[1] * 1e3,
                            "note": "includes the serial float32 mean kernel, scene (re)allocation and two syncs"},
        "from_pinned_host": {"ms_min": t_host[0] * 1e3, "ms_median": t_host[1] * 1e3,
                             "h2d_gbs": n * row_bytes / t_host[0] / 1e9},
        "alg_bytes": alg,
        "cpu_load_ply": {"n": m, "s": t_cpu, "vertices_per_s": m / t_cpu, "kind": "port (util_gau.load_ply, NumPy)"},
        "device_load_ply_from_file": {"n": m, "s": t_file, "vertices_per_s": m / t_file},
        "whole_file": full,
        "peaks": peaks,
    }
    print(json.dumps(line))


if __name__ == "__main__":
    main()
