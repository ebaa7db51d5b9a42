"""Multi-GPU plumbing: one process per GPU (torch.distributed), the frame is partitioned across ranks,
every rank owns a replicated device scene (each builds its own LBVH: the build is deterministic and
takes milliseconds), and the per-rank accumulation buffers are summed with ONE reduce to rank 0
(NCCL over NVLink on GPUs; gloo in the CPU tests).  The path itself has no exchange step - (pixel,
sample) pairs are independent (main.py:186, 199) and the Philox stream is keyed by (pixel, sample,
bounce), so any partition renders exactly the same samples.

Two partitions:
  * tile sharding   - rank g renders ray-index tiles {c : c % G == g} for all samples (config 3);
  * sample sharding - rank g renders the whole frame for samples [g*spp/G, (g+1)*spp/G) (config 5).
"""
import numpy as np


def shard_ray_indices(width, height, shard, n_shards, tile):
    """Host mirror of the device pixel map (ptb_common.cuh PixelMap): ray indices r = ix*H + iy owned
    by `shard`, in pixel-slot order."""
    total = width * height
    if n_shards <= 1:
        return np.arange(total, dtype=np.int64)
    chunks = np.arange(shard, (total + tile - 1) // tile, n_shards, dtype=np.int64)
    idx = (chunks[:, None] * tile + np.arange(tile, dtype=np.int64)[None, :]).reshape(-1)
    return idx[idx < total]


def sample_range(spp, rank, world):
    """Samples [begin, end) of `rank` under sample-range sharding."""
    return (spp * rank) // world, (spp * (rank + 1)) // world


def reduce_frame(full, dst=0, group=None):
    """Sum the full-frame accumulation buffers of all ranks into rank `dst` (single collective)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(full, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return full


def replicate_renderer(scene, device, src=0, group=None, accel=0, make=None):
    """One device scene per rank with the hierarchies built ONCE: rank `src` builds (ptb_scene_create), exports
    both LBVHs into one device buffer (ptb_bvh_export) and broadcasts it (one collective, NCCL over NVLink);
    every other rank imports it (ptb_bvh_import: device-to-device copies instead of sort / tree / refit).
    Small scenes (shared-memory plane list, nothing to export: size 0 is broadcast) and single-process runs
    build locally.  `make(scene, device, accel, bvh)` creates the renderer (tests inject a stand-in)."""
    import torch
    import torch.distributed as dist
    if make is None:
        from .render import Renderer
        make = lambda sc, dev, acc, bvh: Renderer(sc, device=dev, accel=acc, bvh=bvh)
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    if world <= 1:
        return make(scene, device, accel, None)
    rank = dist.get_rank(group)
    r, blob = None, None
    size = torch.zeros(1, dtype=torch.int64, device=device)
    if rank == src:
        r = make(scene, device, accel, None)
        try:
            blob = r.export_bvh()
            size[0] = blob.numel()
        except RuntimeError:        # PtbError "no LBVH": a shared-memory scene, everybody builds its own plane list
            blob = None
    dist.broadcast(size, src=src, group=group)
    n = int(size.item())
    if n == 0:
        return r if r is not None else make(scene, device, accel, None)
    if rank != src:
        blob = torch.empty(n, dtype=torch.uint8, device=device)
    dist.broadcast(blob, src=src, group=group)
    return r if rank == src else make(scene, device, accel, blob)


def render_step(renderer, spp, bounces, seed, width, height, rank=0, world=1, mode="tile", tile=None, full=None,
                flags=0, tonemap=True):
    """One frame on this rank's share of the work; returns the u8 image tensor on rank 0 (else None)
    and the full-frame accumulation buffer."""
    torch = renderer.torch
    tile = tile or height
    if mode == "tile":
        acc, _ = renderer.trace_paths(spp, bounces, seed=seed, width=width, height=height, shard=rank, n_shards=world,
                                      tile=tile, flags=flags)
        if world > 1:
            if full is None:
                full = torch.zeros((width * height, 4), dtype=torch.float32, device=renderer.device)
            else:
                full.zero_()
            px, _ = renderer.pixels(width, height, rank, world, tile)
            renderer.scatter(px, acc, full)
        else:
            full = acc
    elif mode == "sample":
        s0, s1 = sample_range(spp, rank, world)
        if full is None:
            full = torch.zeros((width * height, 4), dtype=torch.float32, device=renderer.device)
        else:
            full.zero_()
        if s1 > s0:
            renderer.trace_paths(s1 - s0, bounces, seed=seed, width=width, height=height, sample_begin=s0, accum=full, flags=flags)
    else:
        raise ValueError(f"unknown sharding mode {mode!r}")
    if world > 1:
        reduce_frame(full, 0)
    img = renderer.image(full, width, height) if (rank == 0 and tonemap) else None
    return img, full
