"""Multi-GPU plumbing for pose batches (the render_all.py / camera_data.csv workload).

The path shards over camera views: every rank holds the whole scene and renders its own slice of the
pose list.  The only collective is ONE broadcast of the scene arrays at load time; nothing is exchanged
per frame.  Works with any torch.distributed backend (NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

import numpy as np

from .util_gau import GaussianData

_FIELDS = ("xyz", "rot", "scale", "opacity", "sh")
_COLS = {"xyz": 3, "rot": 4, "scale": 3, "opacity": 1}


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous, balanced slice [lo, hi) of n_items for `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_round_robin(n_items: int, rank: int, world: int):
    """Interleaved assignment, for trajectories whose per-pose cost varies along the path."""
    return list(range(rank, n_items, world))


def bind_to_gpu_numa_node(device_index: int):
    """Pins the calling process to the CPU cores that are local to the GPU's PCIe root (sysfs `local_cpulist`), BEFORE
    it allocates pinned host buffers: the read-back copies then land in memory of the GPU's own NUMA node.  torchrun
    does not do this.  Returns the core list, or None when sysfs does not tell (then nothing is changed)."""
    import os
    import torch
    try:
        bus = torch.cuda.get_device_properties(device_index).pci_bus_id
        dom = torch.cuda.get_device_properties(device_index).pci_domain_id
        dev = torch.cuda.get_device_properties(device_index).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/local_cpulist"
        with open(path) as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None


def broadcast_scene(gaus, device, src: int = 0):
    """Replicates a scene from rank `src` to every rank.  `gaus` is a GaussianData on the source rank
    (ignored elsewhere).  Returns the five arrays as torch tensors on `device`, in GaussianData order."""
    import torch
    import torch.distributed as dist

    rank = dist.get_rank()
    meta = torch.zeros(2, dtype=torch.int64, device=device)
    if rank == src:
        meta[0], meta[1] = len(gaus), gaus.sh_dim
    dist.broadcast(meta, src=src)
    n, sh_dim = int(meta[0]), int(meta[1])
    out = []
    for f in _FIELDS:
        cols = _COLS.get(f, sh_dim)
        if rank == src:
            t = torch.from_numpy(np.ascontiguousarray(getattr(gaus, f), dtype=np.float32).reshape(n, cols)).to(device)
        else:
            t = torch.empty((n, cols), dtype=torch.float32, device=device)
        dist.broadcast(t, src=src)
        out.append(t)
    return out


def broadcast_scene_resident(renderer, gaus, src: int = 0):
    """Replicates a scene from rank `src` straight into every rank's RESIDENT scene buffers: the source rank
    uploads `gaus` (update_gaussian_data), the others allocate empty buffers (gsr_alloc_gaussians), and the four
    resident planes (xyz, rot, scale+opacity, sh; gsr_scene_buffer) are the broadcast buffers themselves — one
    NCCL broadcast per plane, no second copy of the scene anywhere.  Returns (n, sh_dim)."""
    import torch
    import torch.distributed as dist

    rank = dist.get_rank()
    meta = torch.zeros(2, dtype=torch.int64, device=renderer.device)
    if rank == src:
        meta[0], meta[1] = len(gaus), gaus.sh_dim
        renderer.update_gaussian_data(gaus)
    dist.broadcast(meta, src=src)
    n, sh_dim = int(meta[0]), int(meta[1])
    if rank != src:
        renderer.alloc_gaussians(n, sh_dim)
    if n > 0:
        for t in renderer.scene_buffers():
            dist.broadcast(t, src=src)
    return n, sh_dim


def tensors_to_gaussians(ts) -> GaussianData:
    return GaussianData(*[t.cpu().numpy() for t in ts])
