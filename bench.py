"""bench.py — frames/sec of the forward render path at BASELINE.json's headline configuration.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one frame: one pass of preprocess -> sort -> blend over one camera pose (a new pose every
step; the order is recomputed every frame, nothing is cached between frames).  Prints ONE JSON line.
See DESIGN.md section "Measurement" for what every field means and how the algorithmic bytes are
counted.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[2]: the configuration the metric is quoted on
    "3M_sh3_1080p": dict(n=3_000_000, w=1920, h=1080, mod=3),
    "1M_sh3_1080p": dict(n=1_000_000, w=1920, h=1080, mod=3),   # configs[1]
    "6M_sh3_2160p": dict(n=6_000_000, w=3840, h=2160, mod=3),   # configs[3]
    "small": dict(n=200_000, w=640, h=360, mod=3),              # quick self-test
}
METRIC = "frames/sec @1080p, 3M SH3 Gaussians/B200"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def orbit_pose(k: int, total: int, radius=10.0):
    """Pose k of a closed orbit around the scene at the single-view distance of SURVEY 8(d)."""
    from gaussiansplattingviewer_b200.util import look_at
    ang = 2 * np.pi * (k / max(total, 1))
    eye = np.array([radius * np.sin(ang), 0.0, radius * np.cos(ang)], dtype=np.float32)
    front = -eye / np.linalg.norm(eye)
    return {"camera_front": front.astype(np.float32), "camera_up": np.array([0, -1, 0], np.float32),
            "camera_position": eye, "camera_view": look_at(eye, (0, 0, 0), (0, -1, 0))}


def stage_bytes(n, v, d, tiles, w, h, sh_dim, keys64):
    """Algorithmic bytes per frame of each stage (DESIGN.md 'Kernels'): what must cross HBM once."""
    # K1: xyz for all, rot + scale/opacity + SH for the visible ones; out: 3 record planes, depth key, tile rect
    k1 = 12 * n + (16 + 16 + 4 * sh_dim) * v + 48 * v + 4 * n + 8 * n
    tbits = max(1, int(np.ceil(np.log2(tiles))))
    if keys64:
        passes = -(-(32 + tbits) // 8)
        return {"preprocess": k1, "sort_depth": 0, "ranges": 8 * tiles, "duplicate": 8 * n + 4 * v + 12 * d,
                "sort_tiles": (8 + 24 * passes) * d + 8 * d, "blend": 52 * d + 16 * w * h}
    # depth sort over N: first pass (4 in, 8 out) + three passes (8 in, 8 out); its digit histograms are kept by
    # the preprocess kernel, there is no histogram read
    # duplicate: sorted index (4) + tile rect (8) per Gaussian in, (tile, index) per instance out
    # tile sort: the histograms come from k_tile_scan; pass 1 moves pairs (8 in, 8 out), pass 2 reads pairs and
    # writes only the indices (8 in, 4 out); a third pass exists above 65536 tiles
    tile_passes = 1 if tbits <= 8 else (2 if tbits <= 16 else 3)
    return {"preprocess": k1, "sort_depth": (12 + 16 * 3) * n, "ranges": 8 * tiles, "duplicate": 12 * n + 8 * d,
            "sort_tiles": (16 * tile_passes - 4) * d, "blend": 52 * d + 16 * w * h}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gaussiansplattingviewer_b200 import scenes
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
    from gaussiansplattingviewer_b200.util import Camera

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = WORKLOADS[args.workload]
    n, W, H, mod = wl["n"], wl["w"], wl["h"], wl["mod"]
    K, Wm = args.steps, args.warmup
    r = B200Renderer(W, H, device=local, keys64=args.keys64)
    cam = Camera(H, W)
    # ---- scene: generated on rank 0, replicated ONCE by NCCL broadcast (no collective per frame)
    t_bcast_ms = 0.0
    if rank == 0:
        g = scenes.synthetic_gaussians(n, seed=0)
    if world == 1:
        r.update_gaussian_data(g)
    else:
        shapes = [(n, 3), (n, 4), (n, 3), (n, 1), (n, 48)]
        if rank == 0:
            ts = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (g.xyz, g.rot, g.scale, g.opacity, g.sh)]
        else:
            ts = [torch.empty(s, dtype=torch.float32, device="cuda") for s in shapes]
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for t in ts:
            dist.broadcast(t, src=0)
        e1.record(); torch.cuda.synchronize()
        t_bcast_ms = e0.elapsed_time(e1)
        r.update_gaussian_data_device(*ts)
        del ts
    r.set_scale_modifier(1.0)
    r.set_render_mod(mod)
    r.update_camera_intrin(cam)
    total_poses = world * (K + Wm)

    def pose_of(step):  # every rank walks its own contiguous arc of the orbit
        return orbit_pose(rank * (K + Wm) + step, total_poses)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- region A: device-resident throughput (the `value`)
    for s in range(Wm):
        r.update_camera_pose(cam, True, pose_of(s)); r.draw()
    launches0 = r.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(K):
        r.update_camera_pose(cam, True, pose_of(Wm + s)); r.draw()
    e1.record()
    sync_all()
    ms_a = max_over_ranks(e0.elapsed_time(e1))
    launches = r.launch_count() - launches0
    st = r.stats()
    assert not st.overflow, "instance buffer overflow in the timed region"
    # ---- region B: end to end through the renderer interface, host buffers: pose in, RGB8 image out
    # (the capture loop of main.py:793-917: every pose is rendered and read back; the copy of frame k to
    # pinned host memory overlaps the rendering of frame k+1, every copy completes inside the timed region)
    for s in range(min(Wm, 3)):
        r.update_camera_pose(cam, True, pose_of(s)); r.read_pixels_rgb8_async()
    r.sync_host()
    sync_all()
    t_b0 = time.perf_counter()
    e0.record()
    for s in range(K):
        r.update_camera_pose(cam, True, pose_of(Wm + s)); img8 = r.read_pixels_rgb8_async()
    r.sync_host()
    e1.record()
    sync_all()
    wall_b = (time.perf_counter() - t_b0) * 1e3
    # the copies run on the library's own stream: take the larger of the event time and the host clock
    ms_b = max_over_ranks(max(e0.elapsed_time(e1), wall_b))
    clocks = sampler.stop() if rank == 0 else None
    assert img8.shape == (H, W, 3) and img8.any()
    out = {"ms_a": ms_a, "ms_b": ms_b, "launches": launches, "t_bcast_ms": t_bcast_ms, "clocks": clocks}
    # ---- region C (rank 0): per-stage CUDA-event times over K frames, pair counts, micro-benchmark peaks
    if rank == 0:
        acc = np.zeros(7)
        for s in range(K):
            r.update_camera_pose(cam, True, pose_of(Wm + s))
            _, ss = r.draw_timed()
            acc += [ss.ms_preprocess, ss.ms_sort_depth, ss.ms_ranges, ss.ms_duplicate, ss.ms_sort_tiles, ss.ms_blend,
                    ss.ms_total]
        out["stage_ms"] = dict(zip(["preprocess", "sort_depth", "ranges", "duplicate", "sort_tiles", "blend", "total"],
                                   (acc / K).tolist()))
        out["stats"] = dict(n=n, visible=int(ss.n_visible), instances=int(ss.n_instances), tiles=ss.tiles_x * ss.tiles_y)
        # (pixel, splat) pairs of SURVEY 8(d): per pixel, its position in the 16x16-TILE list when it terminates.
        # The frames above use 32x32 tiles (longer lists per pixel), so the algorithmic figure is counted on a
        # 16x16 re-render of the same pose; it does not depend on how the product cuts the list up.
        from gaussiansplattingviewer_b200 import _gsr
        out["pairs_own_tiling"] = r.count_pairs()
        r.tune(_gsr.GSR_TUNE_TILE, 16)
        r.draw()
        torch.cuda.synchronize()
        out["pairs"] = r.count_pairs()
        r.tune(_gsr.GSR_TUNE_TILE, 32)
        out["peaks"] = r.measure_peaks()
    out.update(rank=rank, world=world, W=W, H=H, n=n, mod=mod, K=K, Wm=Wm, renderer=r, scene=g if rank == 0 else None,
               cam=cam, pose=pose_of(Wm + K - 1))
    if world > 1:
        dist.barrier()
    return out


def sort_gaussian_cpu(xyz, view_mat):
    """The reference's CPU sort, restated line for line (/root/reference/renderer_ogl.py:10-19)."""
    xyz_view = view_mat[None, :3, :3] @ xyz[..., None] + view_mat[None, :3, 3, None]
    depth = xyz_view[:, 2, 0]
    index = np.argsort(depth)
    return index.astype(np.int32).reshape(-1, 1)


def cpu_frame(g, cam, pose, W, H, mod, use_ref):
    """One frame on the host cores: numpy depth sort + CPU rasteriser (the reference's CPU path).
    use_ref: the reference's shaders compiled as C++ (oracle/_ref); else the C port (oracle/gs_oracle.c)."""
    from oracle import oracle as O
    V, P, hff = pose["camera_view"], cam.get_project_matrix(), cam.get_htanfovxy_focal()
    c = O.make_camera(V, P, hff, pose["camera_position"], 1.0, mod, g.sh_dim)
    t0 = time.perf_counter()
    order = sort_gaussian_cpu(g.xyz, V)
    t1 = time.perf_counter()
    if use_ref:
        img = O.ref_render(g, order, c, W, H, clamp_unorm=(mod != -1))
    else:
        img, _, _ = O.render(g, c, W, H, clamp_unorm=(mod != -1))
    t2 = time.perf_counter()
    return img, t1 - t0, t2 - t1


def subsample(g, stride):
    from gaussiansplattingviewer_b200.util_gau import GaussianData
    if stride <= 1:
        return g
    return GaussianData(*[np.ascontiguousarray(a[::stride]) for a in (g.xyz, g.rot, g.scale, g.opacity, g.sh)])


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (numpy sort + its shaders run
    on the host through oracle/_ref, or the C port when _ref is absent), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    from gaussiansplattingviewer_b200 import scenes
    from gaussiansplattingviewer_b200.util import Camera
    from oracle import oracle as O
    wl = WORKLOADS[args.workload]
    n, W, H, mod = wl["n"], wl["w"], wl["h"], wl["mod"]
    K, Wm = args.steps, args.warmup
    use_ref = O.ref_lib() is not None
    # rank 0 is the only rank that works here: take every core this process may run on (torchrun exports
    # OMP_NUM_THREADS=1 for its workers)
    O.set_num_threads(len(os.sched_getaffinity(0)))
    g_full = scenes.synthetic_gaussians(n, seed=0)
    cam = Camera(H, W)
    # bounded sample: every stride-th Gaussian at full resolution, sized from a probe so that
    # (K + W) steps take about args.ref_budget seconds
    probe = subsample(g_full, 64)
    _, ts, tr = cpu_frame(probe, cam, orbit_pose(0, K + Wm), W, H, mod, use_ref)
    est_full = (ts + tr) * 64
    stride = int(max(1, np.ceil(est_full * (K + Wm) / args.ref_budget)))
    g = subsample(g_full, stride)
    for s in range(Wm):
        cpu_frame(g, cam, orbit_pose(s, K + Wm), W, H, mod, use_ref)
    t0 = time.perf_counter()
    for s in range(K):
        cpu_frame(g, cam, orbit_pose(Wm + s, K + Wm), W, H, mod, use_ref)
    dt = time.perf_counter() - t0
    frac = len(g) / n
    fps = frac * K / dt
    sample = (f"every {stride}-th Gaussian ({len(g)} of {n}) at full {W}x{H}; a step = numpy depth sort + CPU "
              f"rasteriser of that sample; frames/s scaled by {frac:.4f}")
    return {"metric": METRIC, "value": fps, "unit": "frames/s", "impl": "reference", "n_gpus": args.gpus, "steps": K,
            "warmup": Wm, "ms_per_step": 1e3 * dt / K / frac, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "n_gaussians": n, "width": W, "height": H, "sh_degree": 3,
                       "render_mod": mod},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": O.num_threads(),
                             "kind": "reference" if use_ref else "port", "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="3M_sh3_1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--keys64", action="store_true", help="one 64-bit-key sort instead of the two-level sort")
    ap.add_argument("--ref-budget", type=float, default=120.0, help="seconds of CPU work for --impl reference")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    # stdout carries exactly one JSON line: everything libraries print there (NCCL's version banner, ...) is
    # sent to stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    if args.impl == "reference":
        line = run_reference(args)
        if line is not None:
            emit(line)
        return
    o = run_ours(args)
    if o["world"] > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    if o["rank"] != 0:
        return
    K, world, W, H, n = o["K"], o["world"], o["W"], o["H"], o["n"]
    hbm_peak, peak_src = load_peaks()
    st, sm = o["stats"], o["stage_ms"]
    sb = stage_bytes(n, st["visible"], st["instances"], st["tiles"], W, H, 48, args.keys64)
    fp32_peak, ex2_peak, copy_gbs = o["peaks"]
    visited, contributed = o["pairs"]
    stages = {}
    for k, b in sb.items():
        ms = sm[k]
        stages[k] = {"ms": round(ms, 4), "alg_bytes": int(b), "gbs": round(b / (ms * 1e-3) / 1e9, 1) if ms > 0 else None,
                     "frac_hbm": round(b / (ms * 1e-3) / 1e9 / hbm_peak, 4) if ms > 0 else None}
    blend_flops = 22.0 * visited
    stages["blend"].update({"pairs_visited": int(visited), "pairs_contributed": int(contributed),
                            "pairs_visited_32x32_lists": int(o["pairs_own_tiling"][0]),
                            "fp32_tflops": round(blend_flops / (sm["blend"] * 1e-3) / 1e12, 2),
                            "frac_fp32": round(blend_flops / (sm["blend"] * 1e-3) / 1e12 / fp32_peak, 4),
                            "ex2_tops": round(contributed / (sm["blend"] * 1e-3) / 1e12, 3),
                            "frac_ex2": round(contributed / (sm["blend"] * 1e-3) / 1e12 / ex2_peak, 4)})
    # DRAM bytes per launch from the committed ncu --set full capture of the same workload (not measured in this run)
    traffic = {}
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tj = json.load(f)
        if tj.get("workload") == args.workload and not args.keys64:
            traffic = tj.get("per_stage", {})
    except Exception:
        traffic = {}
    for k in stages:
        stages[k]["ncu_dram_bytes"] = traffic.get(k)
    dom = max(sb, key=lambda k: sm[k])
    if dom == "blend":
        roof = {"kernel": "k_blend", "bound": "fp32", "achieved": stages["blend"]["fp32_tflops"], "peak": round(fp32_peak, 2),
                "unit": "TFLOP/s", "frac": stages["blend"]["frac_fp32"], "traffic": traffic.get("blend"),
                "peak_source": "FFMA micro-benchmark in this run (gsr_measure_peaks)",
                "alg_flops_per_launch": blend_flops, "hbm_gbs": stages["blend"]["gbs"], "hbm_frac": stages["blend"]["frac_hbm"]}
    else:
        roof = {"kernel": dom, "bound": "hbm", "achieved": stages[dom]["gbs"], "peak": hbm_peak, "unit": "GB/s",
                "frac": stages[dom]["frac_hbm"], "traffic": traffic.get(dom), "peak_source": peak_src,
                "alg_bytes_per_launch": stages[dom]["alg_bytes"]}
    fps = world * K / (o["ms_a"] * 1e-3)
    e2e_fps = world * K / (o["ms_b"] * 1e-3)
    line = {"metric": METRIC, "value": round(fps, 2), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": o["Wm"],
            "ms_per_step": round(o["ms_a"] / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "n_gaussians": n, "width": W, "height": H, "sh_degree": 3,
                       "render_mod": o["mod"], "sort": "keys64" if args.keys64 else "two-level (depth over N, tile over D)",
                       "order_recomputed_every_frame": True, "poses": "orbit r=10 about the origin, one new pose per step",
                       "l2": "scene buffers (%.0f MB) exceed the 126 MB L2" % (n * 236 / 1e6),
                       "visible": st["visible"], "instances": st["instances"],
                       "mean_tile_list": round(st["instances"] / st["tiles"], 1), "tile_edge_px": 32},
            "clocks": o["clocks"],
            "e2e": {"value": round(e2e_fps, 2), "unit": "frames/s", "h2d_bytes_per_step": 16 * 4 * 2 + 9 * 4 + 6 * 4,
                    "d2h_bytes_per_step": W * H * 3,
                    "api": "B200Renderer.update_camera_pose + read_pixels_rgb8_async (gsr_render_rgb8_host_async) per step, "
                           "sync_host at the end: pose from host, RGB8 frame to pinned host memory; the copy of frame k "
                           "overlaps the rendering of frame k+1, all copies finish inside the timed region"},
            "gpu_launches": int(o["launches"]), "roofline": roof, "stages": stages,
            "peaks": {"hbm_gbs": hbm_peak, "hbm_source": peak_src, "fp32_tflops": round(fp32_peak, 2),
                      "ex2_tops": round(ex2_peak, 3), "copy_gbs_this_run": round(copy_gbs, 1)},
            "scene_broadcast_ms": round(o["t_bcast_ms"], 3)}
    # BASELINE.json's metric also names "sort Gkeys/s": the two sorts inside the frame (keys through a complete
    # key/value sort per second: 32 depth bits over N, the tile bits over D) and, at N=1, the standalone
    # gsr_sort_pairs on 2^24 random 32-bit keys with 32-bit values (all 32 bits, CUDA events, best of 5)
    line["sort"] = {"depth_sort_gkeys_s": round(n / (sm["sort_depth"] * 1e-3) / 1e9, 2) if sm["sort_depth"] > 0 else None,
                    "tile_sort_gkeys_s": round(st["instances"] / (sm["sort_tiles"] * 1e-3) / 1e9, 2) if sm["sort_tiles"] > 0 else None}
    if world == 1:
        import torch
        m = 1 << 24
        gen = torch.Generator(device="cuda").manual_seed(1)
        best = None
        for _ in range(5):
            tk = torch.randint(-2 ** 31, 2 ** 31 - 1, (m,), dtype=torch.int32, device="cuda", generator=gen)
            tv = torch.arange(m, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            o["renderer"].sort_pairs(tk, tv)
            e1.record()
            torch.cuda.synchronize()
            t = e0.elapsed_time(e1)
            best = t if best is None else min(best, t)
        line["sort"]["pairs_2p24_u32_gkeys_s"] = round(m / (best * 1e-3) / 1e9, 2)
        line["sort"]["pairs_2p24_u32_ms"] = round(best, 4)
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as O
        img, t_sort, t_rast = cpu_frame(o["scene"], o["cam"], o["pose"], W, H, o["mod"], use_ref=False)
        gpu_img = o["renderer"].read_pixels_rgba32f()
        line["cpu_baseline"] = {"value": round(1.0 / (t_sort + t_rast), 4), "unit": "frames/s", "cores": O.num_threads(),
                                "kind": "port", "sample": "one full frame of the same workload (last timed pose): numpy "
                                "depth sort %.2f s + oracle/gs_oracle.c rasteriser %.2f s" % (t_sort, t_rast),
                                "max_abs_diff_vs_gpu": float(np.abs(gpu_img - img).max())}
    emit(line)


if __name__ == "__main__":
    main()
