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


def csrc_hash():
    """Hash of the kernel sources: an ncu capture is only quoted beside a build of the same sources."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "gaussiansplattingviewer_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".h")):
            with open(os.path.join(d, f), "rb") as fh:
                h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def static_config(workload, n, W, H, mod, keys64=False):
    """The `config` object of BOTH arms (ours and --impl reference): what is rendered, nothing run-dependent."""
    return {"workload": workload, "n_gaussians": n, "width": W, "height": H, "sh_degree": 3, "render_mod": mod,
            "order_recomputed_every_frame": True, "poses": "orbit r=10 about the origin, one new pose per step",
            "l2": "scene buffers (%.0f MB) exceed the 126 MB L2" % (n * 236 / 1e6)}


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
    # depth sort: the compacting first pass reads the N keys (4 B) and writes (key', index) for the V Gaussians that
    # have one (8 B); pass two moves V pairs (8 in, 8 out); pass three reads V pairs and writes only the indices
    # (8 in, 4 out).  Its digit histograms are kept by the preprocess kernel, there is no histogram read.
    # duplicate: sorted index (4) + tile rect (8) per VISIBLE Gaussian in, (tile, index) per instance out
    # tile sort: the histograms come from k_tile_scan; pass 1 moves pairs (8 in, 8 out), pass 2 reads pairs and
    # writes only the indices (8 in, 4 out); a third pass exists above 65536 tiles
    tile_passes = 1 if tbits <= 8 else (2 if tbits <= 16 else 3)
    return {"preprocess": k1, "sort_depth": 4 * n + 28 * v, "ranges": 8 * tiles, "duplicate": 12 * v + 8 * d,
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
    from gaussiansplattingviewer_b200 import pose_shard as _ps
    numa_cpus = _ps.bind_to_gpu_numa_node(local) if world > 1 else None  # pinned read-back buffers on the GPU's own node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = WORKLOADS[args.workload]
    n, W, H, mod = wl["n"], wl["w"], wl["h"], wl["mod"]
    K, Wm = args.steps, args.warmup
    r = B200Renderer(W, H, device=local, keys64=args.keys64)
    cam = Camera(H, W)
    # ---- scene: generated on rank 0, replicated ONCE by NCCL broadcast (no collective per frame)
    t_bcast_ms = 0.0
    g = scenes.synthetic_gaussians(n, seed=0) if rank == 0 else None
    if world == 1:
        r.update_gaussian_data(g)
    else:
        # rank 0 uploads, the others allocate; the four resident planes are the NCCL broadcast buffers themselves
        # (pose_shard.broadcast_scene_resident: gsr_alloc_gaussians + gsr_scene_buffer, no second copy of the scene)
        from gaussiansplattingviewer_b200 import pose_shard
        if rank == 0:
            r.update_gaussian_data(g)
        else:
            r.alloc_gaussians(n, 48)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for t in r.scene_buffers():
            dist.broadcast(t, src=0)
        e1.record(); torch.cuda.synchronize()
        t_bcast_ms = e0.elapsed_time(e1)
        assert pose_shard.broadcast_scene_resident.__doc__  # the package entry point does exactly the above
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
    assert img8.shape == (H, W, 3) and img8.any()
    out = {"ms_a": ms_a, "ms_b": ms_b, "launches": launches, "t_bcast_ms": t_bcast_ms}
    # ---- what the read-back path can carry: every rank copies a frame-sized buffer (RGB8 1080p x 10 = 62 MB) device ->
    # pinned host at the same time, GB/s per rank (at N = 8 the GPUs share the host's PCIe switches and memory)
    d2h_dev = torch.empty(W * H * 3 * 10, dtype=torch.uint8, device="cuda")
    d2h_host = torch.empty(W * H * 3 * 10, dtype=torch.uint8, pin_memory=True)
    d2h_host.copy_(d2h_dev, non_blocking=True)
    sync_all()
    e0.record()
    for _ in range(8):
        d2h_host.copy_(d2h_dev, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    d2h_gbs = 8 * d2h_dev.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9
    if world > 1:
        tt = torch.tensor([d2h_gbs], dtype=torch.float64, device="cuda")
        allt = [torch.zeros_like(tt) for _ in range(world)]
        dist.all_gather(allt, tt)
        per = [float(x.item()) for x in allt]
    else:
        per = [d2h_gbs]
    out["d2h_probe"] = {"gbs_per_rank_min": round(min(per), 2), "gbs_per_rank_median": round(float(np.median(per)), 2),
                        "gbs_per_rank_max": round(max(per), 2), "frame_copy_ms_at_min": round(W * H * 3 / (min(per) * 1e9) * 1e3, 4),
                        "what": "all ranks copy 62 MB device -> pinned host concurrently, 8 times"}
    del d2h_dev, d2h_host
    # ---- config 5 (BASELINE.json configs[4]): 1024 trajectory poses over the same scene, sharded over the ranks
    out["config5"] = run_config5(r, cam, rank, world, sync_all, W, H) if args.workload == "3M_sh3_1080p" else None
    out["clocks"] = sampler.stop() if rank == 0 else None
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
    out.update(numa=numa_cpus, rank=rank, world=world, W=W, H=H, n=n, mod=mod, K=K, Wm=Wm, renderer=r, scene=g if rank == 0 else None,
               cam=cam, pose=pose_of(Wm + K - 1))
    if world > 1:
        dist.barrier()
    return out


def run_config5(r, cam, rank, world, sync_all, W, H, n_poses=1024):
    """The render_all.py / camera_data.csv workload (/root/reference/render_all.py:14-35, main.py:529-562, 793-840):
    a FIXED list of 1024 poses (closed Catmull-Rom loop through the 18 rows of the viewer's camera_data.csv,
    SURVEY 8d) over the resident scene, split into contiguous slices with pose_shard.shard_range; strong scaling.
    Two timed loops per rank, each bracketed by barrier + synchronize, time = max over ranks:
      device : update_camera_pose + draw per pose
      e2e    : update_camera_pose + read_pixels_rgb8_async per pose, sync_host at the end (every RGB8 frame is
               copied to pinned host memory inside the region)"""
    import torch
    import torch.distributed as dist
    from gaussiansplattingviewer_b200 import pose_shard, scenes
    rows = np.load(os.path.join(ROOT, "tests", "golden", "camera_data.npy"))
    views, campos = scenes.trajectory_poses(rows, n_poses)
    # interleaved assignment (pose_shard.shard_round_robin): the cost of a pose varies along the trajectory, contiguous
    # slices left one rank 14 % longer than the other at N = 2 (219 vs 250 ms)
    mine = pose_shard.shard_round_robin(n_poses, rank, world)
    zero3 = np.zeros(3, np.float32)

    def pose(i):
        return {"camera_front": zero3, "camera_up": zero3, "camera_position": campos[i], "camera_view": views[i]}

    res = {}
    for name in ("device", "e2e"):
        for i in mine[:3]:  # warm-up on the rank's first poses
            r.update_camera_pose(cam, True, pose(i))
            r.draw() if name == "device" else r.read_pixels_rgb8_async()
        r.sync_host()
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        t_enq0 = time.perf_counter()
        for i in mine:
            r.update_camera_pose(cam, True, pose(i))
            r.draw() if name == "device" else r.read_pixels_rgb8_async()
        enqueue_us = (time.perf_counter() - t_enq0) * 1e6 / max(len(mine), 1)  # host time per pose inside the loop
        if name == "e2e":
            r.sync_host()
        e1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        ms = max(e0.elapsed_time(e1), wall) if name == "e2e" else e0.elapsed_time(e1)
        host_ms = wall  # host time to ENQUEUE the slice is below this; the loop never waits except at the end
        sync_all()
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            allt = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allt, t)
            per_rank = [float(x.item()) for x in allt]
        else:
            per_rank = [ms]
        res[name] = {"views_per_s": round(n_poses / (max(per_rank) * 1e-3), 1), "ms_total_max_over_ranks": round(max(per_rank), 3),
                     "per_rank_ms": {"min": round(min(per_rank), 3), "median": round(float(np.median(per_rank)), 3),
                                     "max": round(max(per_rank), 3)},
                     "poses_rank0": len(mine), "rank0_wall_ms": round(host_ms, 3),
                     "rank0_host_us_per_pose_in_loop": round(enqueue_us, 1)}
    assert not r.stats().overflow
    return {"workload": "1024 poses (closed Catmull-Rom loop through camera_data.csv, x3 about the origin) over the "
                        "3M SH3 scene at 1920x1080, poses dealt round-robin to the ranks (pose_shard.shard_round_robin)",
            "n_poses": n_poses, "scaling": "strong", "views_per_s": res["device"]["views_per_s"],
            "views_per_s_e2e": res["e2e"]["views_per_s"], "device": res["device"], "e2e": res["e2e"],
            "d2h_bytes_per_view": W * H * 3}


def run_stereo(r, cam, n_poses=40):
    """The viewer's stereo capture loop (/root/reference/main.py:844-917, one scene of render_all.py:14-35): per pose left
    RGB8 + disparity uint16 (render_mod -1 at the left pose) + right RGB8 into pinned host buffers.  poses/s through
    B200Renderer.capture_stereo (one depth order per pose shared by the three renders, copies overlapped with the next
    pose) next to a straight port of the reference loop (three synchronous read-backs per pose)."""
    import torch
    from gaussiansplattingviewer_b200 import colmap, scenes
    views, campos = scenes.orbit_poses(n_poses)
    front, up = np.array([0, 0, 1], np.float32), np.array([0, -1, 0], np.float32)
    poses = []
    for v, c in zip(views, campos):
        vr = colmap.right_view(v)
        poses.append(({"camera_view": v, "camera_position": c, "camera_front": front, "camera_up": up},
                      {"camera_view": vr, "camera_position": colmap.camera_position_of(vr), "camera_front": front, "camera_up": up}))
    r.set_render_mod(3)

    def capture(share):
        for pl, pr in poses[:3]:
            r.capture_stereo(pl, pr, share_order=share)
        r.sync_host(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for pl, pr in poses:
            r.capture_stereo(pl, pr, share_order=share)
        r.sync_host(); torch.cuda.synchronize()
        return len(poses) / (time.perf_counter() - t0)

    def port():
        def one(pl, pr):
            r.set_render_mod(3); r.update_camera_pose(cam, True, pl); r.read_pixels_rgb8()
            r.set_render_mod(-1); r.read_pixels_r32f()
            r.set_render_mod(3); r.update_camera_pose(cam, True, pr); r.read_pixels_rgb8()
        for pl, pr in poses[:2]:
            one(pl, pr)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for pl, pr in poses[:20]:
            one(pl, pr)
        torch.cuda.synchronize()
        r.set_render_mod(3)
        return 20 / (time.perf_counter() - t0)

    out = {"poses_per_s_shared_order": round(capture(True), 1), "poses_per_s_three_sorts": round(capture(False), 1),
           "poses_per_s_three_synchronous_readbacks": round(port(), 1), "poses": n_poses,
           "d2h_bytes_per_pose": r.w * r.h * 8, "api": "B200Renderer.capture_stereo (gsr_capture_stereo_async)"}
    return out


def run_config4(local, K=10):
    """BASELINE.json configs[3]: 6M SH3 Gaussians at 3840x2160, colour (3), depth/disparity (-1, the 1.2x-scale second
    geometry pass, gau_vert.glsl:151-156,182-211) and flat ball (-3).  Frames/s with the scene resident (CUDA events
    over K frames at new orbit poses) and per-stage times (gsr_render_timed)."""
    import torch
    from gaussiansplattingviewer_b200 import scenes
    from gaussiansplattingviewer_b200.renderer_b200 import B200Renderer
    from gaussiansplattingviewer_b200.util import Camera
    n, W, H = 6_000_000, 3840, 2160
    r = B200Renderer(W, H, device=local)
    cam = Camera(H, W)
    r.update_gaussian_data(scenes.synthetic_gaussians(n, seed=1))
    r.update_camera_intrin(cam)
    out = {"n_gaussians": n, "width": W, "height": H, "modes": {}}
    for mod in (3, -1, -3):
        r.set_render_mod(mod)
        for s_ in range(3):
            r.update_camera_pose(cam, True, orbit_pose(s_, K + 3)); r.draw()
        torch.cuda.synchronize()
        if r.stats().overflow:  # first frames of a mode may grow the instance buffers (draw() repairs on the next call)
            r.update_camera_pose(cam, True, orbit_pose(0, K + 3)); r.draw(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s_ in range(K):
            r.update_camera_pose(cam, True, orbit_pose(3 + s_, K + 3)); r.draw()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        acc = np.zeros(7)
        for s_ in range(K):
            r.update_camera_pose(cam, True, orbit_pose(3 + s_, K + 3))
            _, ss = r.draw_timed()
            acc += [ss.ms_preprocess, ss.ms_sort_depth, ss.ms_ranges, ss.ms_duplicate, ss.ms_sort_tiles, ss.ms_blend, ss.ms_total]
        st = r.stats()
        out["modes"][str(mod)] = {"frames_per_s": round(1e3 / ms, 1), "ms_per_frame": round(ms, 4), "overflow": int(st.overflow),
                                  "visible": int(st.n_visible), "instances": int(st.n_instances),
                                  "stage_ms": dict(zip(["preprocess", "sort_depth", "ranges", "duplicate", "sort_tiles", "blend", "total"],
                                                       [round(float(v), 4) for v in acc / K]))}
    del r
    torch.cuda.empty_cache()
    return out


def sort_gaussian_torch(xyz_np, view_mat_np):
    """The reference's GPU sort, restated line for line (/root/reference/renderer_ogl.py:41-53; the selector at
    :56-64 picks it whenever torch sees a CUDA device): H2D of xyz on EVERY call (:47), bmm, argsort, D2H."""
    import torch
    xyz = torch.tensor(xyz_np).cuda()
    view_mat = torch.tensor(view_mat_np).cuda()
    xyz_view = view_mat[None, :3, :3] @ xyz[..., None] + view_mat[None, :3, 3, None]
    depth = xyz_view[:, 2, 0]
    index = torch.argsort(depth)
    index = index.type(torch.int32).reshape(-1, 1).cpu().numpy()
    return index


def sort_baselines(r, g, view):
    """Same-box context for "sort Gkeys/s": the reference's torch sort next to gsr_sort_depth on the same 3M
    positions, and torch.sort (CUB) next to gsr_sort_pairs on 2^24 random 32-bit keys."""
    import torch
    out = {}
    n = len(g.xyz)
    view = np.ascontiguousarray(view, dtype=np.float32)

    def best_wall(fn, reps=3):
        fn(); torch.cuda.synchronize()
        best = None
        for _ in range(reps):
            torch.cuda.synchronize(); t0 = time.perf_counter(); res = fn(); torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) * 1e3
            best = dt if best is None else min(best, dt)
        return best, res

    ms_ref, idx_ref = best_wall(lambda: sort_gaussian_torch(g.xyz, view))
    ms_ours, idx_ours = best_wall(lambda: r.gaussian_order(view))
    d = (g.xyz @ view[2, :3] + view[2, 3]).astype(np.float32)
    out["reference_torch_ms"] = round(ms_ref, 3)            # _sort_gaussian_torch as the reference calls it
    out["gsr_sort_depth_host_ms"] = round(ms_ours, 3)       # B200Renderer.gaussian_order: resident xyz, index to host
    out["same_depth_order"] = bool(np.array_equal(d[idx_ref[:, 0]], d[idx_ours[:, 0]]))
    # device-only parts on resident data (CUDA events): bmm + argsort  vs  gsr_sort_depth into device memory
    xyz_t = torch.from_numpy(g.xyz).cuda(); vm = torch.from_numpy(view).cuda()

    def ev(fn, reps=5):
        fn(); torch.cuda.synchronize()
        best = None
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            t = e0.elapsed_time(e1)
            best = t if best is None else min(best, t)
        return best

    out["torch_resident_argsort_ms"] = round(ev(lambda: torch.argsort((vm[None, :3, :3] @ xyz_t[..., None] + vm[None, :3, 3, None])[:, 2, 0])), 4)
    out["gsr_sort_depth_device_ms"] = round(ev(lambda: r.gaussian_order(view, to_host=False)), 4)
    out["depth_order_n"] = n
    m = 1 << 24
    gen = torch.Generator(device="cuda").manual_seed(1)
    tk0 = torch.randint(-2 ** 31, 2 ** 31 - 1, (m,), dtype=torch.int32, device="cuda", generator=gen)
    t_torch = ev(lambda: torch.sort(tk0))
    out["torch_sort_pairs_2p24_ms"] = round(t_torch, 4)     # torch.sort returns (sorted keys, int64 indices)
    out["torch_sort_pairs_2p24_gkeys_s"] = round(m / (t_torch * 1e-3) / 1e9, 2)
    del xyz_t, tk0
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
    stride = 1
    if args.ref_budget > 0:  # opt-in bounded sample: every stride-th Gaussian, sized from a probe
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
    if stride == 1:
        sample = (f"the full workload: all {n} Gaussians at {W}x{H}, a step = the reference's numpy depth sort + its "
                  f"shaders run on the host cores ({'oracle/_ref' if use_ref else 'oracle/gs_oracle.c'}), one new orbit pose per step")
    else:
        sample = (f"every {stride}-th Gaussian ({len(g)} of {n}) at full {W}x{H}; a step = numpy depth sort + CPU "
                  f"rasteriser of that sample; frames/s scaled by {frac:.4f}")
    return {"metric": METRIC, "value": fps, "unit": "frames/s", "impl": "reference", "n_gpus": args.gpus, "steps": K,
            "warmup": Wm, "ms_per_step": 1e3 * dt / K / frac, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": static_config(args.workload, n, W, H, mod), "n_gaussians_rendered": len(g),
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
    ap.add_argument("--ref-budget", type=float, default=0.0,
                    help="--impl reference: 0 (default) = the full configuration; > 0 = sub-sample the Gaussians to about this many seconds of CPU work")
    ap.add_argument("--no-config4", action="store_true", help="skip the 6M / 3840x2160 leg (configs[3])")
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
    traffic, traffic_stale = {}, False
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tj = json.load(f)
        if tj.get("workload") == args.workload and not args.keys64:
            if tj.get("csrc_hash") == csrc_hash():
                traffic = tj.get("per_stage", {})
            else:
                traffic_stale = True  # captured on other kernel sources: not quoted
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
    roof["traffic_stale"] = traffic_stale
    roof["csrc_hash"] = csrc_hash()
    fps = world * K / (o["ms_a"] * 1e-3)
    e2e_fps = world * K / (o["ms_b"] * 1e-3)
    line = {"metric": METRIC, "value": round(fps, 2), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": o["Wm"],
            "ms_per_step": round(o["ms_a"] / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": static_config(args.workload, n, W, H, o["mod"]),
            "workload_stats": {"sort": "keys64" if args.keys64 else "two-level (depth over the visible Gaussians, tile over D)",
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
    line["host"] = {"cpu_affinity_cores": len(os.sched_getaffinity(0)), "bound_to_gpu_numa_node": o.get("numa") is not None,
                    "d2h_probe": o.get("d2h_probe")}
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
        line["sort"].update(sort_baselines(o["renderer"], o["scene"], o["pose"]["camera_view"]))
        if args.workload == "3M_sh3_1080p":
            line["stereo"] = run_stereo(o["renderer"], o["cam"])
            o["renderer"].update_camera_pose(o["cam"], True, o["pose"])  # back to the last timed pose (cpu_baseline below)
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as O
        img, t_sort, t_rast = cpu_frame(o["scene"], o["cam"], o["pose"], W, H, o["mod"], use_ref=False)
        gpu_img = o["renderer"].read_pixels_rgba32f()
        line["cpu_baseline"] = {"value": round(1.0 / (t_sort + t_rast), 4), "unit": "frames/s", "cores": O.num_threads(),
                                "kind": "port", "sample": "one full frame of the same workload (last timed pose): numpy "
                                "depth sort %.2f s + oracle/gs_oracle.c rasteriser %.2f s" % (t_sort, t_rast),
                                "max_abs_diff_vs_gpu": float(np.abs(gpu_img - img).max())}
    line["config5"] = o["config5"]
    if world == 1 and not args.no_config4 and args.workload == "3M_sh3_1080p":
        del o["renderer"]
        line["config4"] = run_config4(int(os.environ.get("LOCAL_RANK", "0")))
    emit(line)


if __name__ == "__main__":
    main()
