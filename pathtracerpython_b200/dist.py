"""Multi-GPU plumbing: one process per GPU (torch.distributed), the frame is partitioned across ranks,
every rank owns a replicated device scene (each builds its own LBVH - the build is deterministic and
takes milliseconds - or receives rank 0's, replicate_renderer), and the frame is assembled on rank 0 with ONE
collective (NCCL over NVLink on GPUs; gloo in the CPU tests): a gather of the owned shards for tile sharding,
a sum-reduce of the partial frames for sample-range sharding.  The path itself has no exchange step - (pixel,
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
    """Sum the full-frame accumulation buffers of all ranks into rank `dst` (single collective).  Used by
    sample-range sharding, where every rank holds a partial sum of EVERY pixel."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(full, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return full


def shard_pixel_count(width, height, shard, n_shards, tile):
    """Number of pixels `shard` owns (host mirror of ptb_pixel_count)."""
    total = width * height
    if n_shards <= 1:
        return total
    stride = tile * n_shards
    full, rem = divmod(total, stride)
    lo = shard * tile
    return full * tile + (min(rem - lo, tile) if rem > lo else 0)


def gather_frame(shard_accum, width, height, rank, world, tile, full, scatter, dst=0, group=None):
    """Tile sharding: every pixel is owned by exactly one rank, so nothing needs summing - the owned shards (1/world
    of the frame each) travel to rank `dst` with ONE gather and are scattered into the full frame there:
    (world - 1) / world of a frame crosses NVLink instead of world - 1 full frames into a reduce.
    `scatter(g, shard, full)` places shard g's pixel slots at their ray indices (Renderer.scatter on the GPU).
    Shards are padded to the largest one because gather wants equal sizes."""
    import torch
    import torch.distributed as dist
    n_max = max(shard_pixel_count(width, height, g, world, tile) for g in range(world))
    if shard_accum.shape[0] != n_max:
        padded = torch.zeros((n_max, 4), dtype=shard_accum.dtype, device=shard_accum.device)
        padded[:shard_accum.shape[0]] = shard_accum
    else:
        padded = shard_accum.contiguous()
    parts = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
    dist.gather(padded, gather_list=parts, dst=dst, group=group)
    if rank == dst:
        for g in range(world):
            scatter(g, parts[g][:shard_pixel_count(width, height, g, world, tile)], full)
    return full


def collective_info(width, height, world, mode):
    """What crosses NVLink per frame (reported by bench.py)."""
    frame = width * height * 16
    if world <= 1:
        return {"op": None, "bytes": 0}
    if mode == "tile":
        return {"op": "gather", "bytes": frame * (world - 1) // world, "note": "owned shards to rank 0, float4 per pixel"}
    return {"op": "reduce", "bytes": frame * (world - 1), "note": "full-frame partial sums to rank 0, float4 per pixel"}


def replicate_renderer(scene, device, src=0, group=None, accel=0, make=None):
    """One device scene per rank with the hierarchies built ONCE: rank `src` builds (ptb_scene_create), exports
    both LBVHs into one device buffer (ptb_bvh_export) and broadcasts it (one collective, NCCL over NVLink);
    every other rank imports it (ptb_bvh_import: device-to-device copies instead of sort / tree / refit).
    Small scenes (shared-memory plane list, nothing to export: size 0 is broadcast) and single-process runs
    build locally.  A failure on `src` (out of memory, CUDA error) is broadcast as size -1 so that every rank
    raises instead of waiting in the collective.  `make(scene, device, accel, bvh)` creates the renderer
    (tests inject a stand-in)."""
    import torch
    import torch.distributed as dist
    if make is None:
        from .render import Renderer
        make = lambda sc, dev, acc, bvh: Renderer(sc, device=dev, accel=acc, bvh=bvh)
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    if world <= 1:
        return make(scene, device, accel, None)
    rank = dist.get_rank(group)
    r, blob, err = None, None, None
    size = torch.zeros(1, dtype=torch.int64, device=device)
    if rank == src:
        try:
            r = make(scene, device, accel, None)
            if r.info()["accel"] == "lbvh":          # asked, not inferred from an exception
                blob = r.export_bvh()
                size[0] = blob.numel()
        except Exception as e:                       # noqa: BLE001 - whatever it is, the other ranks must hear of it
            err = e
            size[0] = -1
    dist.broadcast(size, src=src, group=group)
    n = int(size.item())
    if n < 0:
        if err is not None:
            raise err
        raise RuntimeError(f"replicate_renderer: rank {src} failed to build the scene")
    if n == 0:
        return r if r is not None else make(scene, device, accel, None)
    if rank != src:
        blob = torch.empty(n, dtype=torch.uint8, device=device)
    dist.broadcast(blob, src=src, group=group)
    return r if rank == src else make(scene, device, accel, blob)


def gather_scene_arrays(scene, device, rank=None, world=None, group=None):
    """The big scene arrays (FP64 vertices and normals, object ids: 100 B per triangle, 1.0 GB at 10^7 triangles) on
    every rank's GPU with each byte crossing PCIe ONCE: rank g uploads rows [g n / G, (g + 1) n / G) from its
    (page-locked) host copy and one all-gather over NVLink completes the arrays everywhere.  Returns the tuple
    `Renderer(scene, device_arrays=...)` takes.  Every rank still holds the host scene (the small tables - materials,
    light areas - are read from it); what is saved is (G - 1) / G of the host-to-device traffic per rank."""
    import torch
    import torch.distributed as dist
    if world is None:
        world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        rank = dist.get_rank(group) if world > 1 else 0
    flat = scene.flatten() if hasattr(scene, "flatten") else scene
    n = int(np.asarray(flat["verts"]).shape[0])
    per = (n + world - 1) // world
    lo, hi = min(n, rank * per), min(n, (rank + 1) * per)
    out = []
    for key, dt, width in (("verts", torch.float64, 9), ("normals", torch.float64, 3), ("tri_obj", torch.int32, 1)):
        host = torch.from_numpy(np.asarray(flat[key]).reshape(n, width))
        mine = torch.zeros((per, width), dtype=dt, device=device)
        mine[:hi - lo].copy_(host[lo:hi], non_blocking=True)       # this rank's rows, the only ones it reads from the host
        if world > 1:
            full = torch.empty((per * world, width), dtype=dt, device=device)
            dist.all_gather_into_tensor(full, mine, group=group)
        else:
            full = mine
        out.append(full[:n].contiguous() if full.shape[0] != n else full)
    return out[0].view(n, 3, 3), out[1], out[2].view(n)


def render_step(renderer, spp, bounces, seed, width, height, rank=0, world=1, mode="tile", tile=None, full=None,
                flags=0, tonemap=True):
    """One frame on this rank's share of the work; returns the u8 image tensor on rank 0 (else None)
    and the full-frame accumulation buffer (complete on rank 0)."""
    torch = renderer.torch
    tile = tile or height
    if mode == "tile":
        acc, _ = renderer.trace_paths(spp, bounces, seed=seed, width=width, height=height, shard=rank, n_shards=world,
                                      tile=tile, flags=flags)
        if world > 1:
            if full is None and rank == 0:
                full = torch.empty((width * height, 4), dtype=torch.float32, device=renderer.device)

            def scatter(g, shard, dst_full):
                px, _ = renderer.pixels(width, height, g, world, tile)
                renderer.scatter(px, shard, dst_full)
            gather_frame(acc, width, height, rank, world, tile, full, scatter)
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
        if world > 1:
            reduce_frame(full, 0)
    else:
        raise ValueError(f"unknown sharding mode {mode!r}")
    img = renderer.image(full, width, height) if (rank == 0 and tonemap) else None
    return img, full


def multi_gpu_check(renderer, rank, world, width=256, height=256, spp=8, bounces=3, seed=9, lbvh_scene=None):
    """The N-GPU frame against rank 0's own 1-GPU frame (run by every rank, result on rank 0): tile sharding must be
    BITWISE identical (every pixel is computed by one rank, in the same order); sample-range sharding equal up to
    float32 summation order.  With `lbvh_scene`: a hierarchy built on rank 0 and broadcast renders bit-identically to
    the locally built one on every rank.  -> dict for bench.py's JSON line / tests/multi_gpu_check.py."""
    import torch
    import torch.distributed as dist
    _, full_t = render_step(renderer, spp, bounces, seed, width, height, rank, world, "tile", height, tonemap=False)
    if rank == 0:
        full_t = full_t.clone()
    _, full_s = render_step(renderer, spp, bounces, seed, width, height, rank, world, "sample", tonemap=False)
    out = {"frame": f"{width}x{height}x{spp}spp_b{bounces}", "ranks": world}
    if rank == 0:
        ref, _ = renderer.trace_paths(spp, bounces, seed=seed, width=width, height=height)
        out["tile_bitwise"] = bool(torch.equal(ref, full_t))
        out["sample_max_abs"] = float((ref - full_s).abs().max().item())
        out["sample_max_rel"] = float(((ref - full_s).abs() / ref.abs().clamp_min(1e-6)).max().item())
    if lbvh_scene is not None:
        from .render import Renderer
        rep = replicate_renderer(lbvh_scene, renderer.device.index, src=0, accel=2)
        own = Renderer(lbvh_scene, device=renderer.device.index, accel=2)
        a_rep, _ = rep.trace_paths(2, 3, seed=4, width=96, height=96)
        a_own, _ = own.trace_paths(2, 3, seed=4, width=96, height=96)
        same = torch.tensor([1 if torch.equal(a_rep, a_own) else 0], dtype=torch.int32, device=renderer.device)
        if world > 1:
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
        arr = Renderer(lbvh_scene, device=renderer.device.index, accel=2,
                       device_arrays=gather_scene_arrays(lbvh_scene, renderer.device, rank, world))
        a_arr, _ = arr.trace_paths(2, 3, seed=4, width=96, height=96)
        same_arr = torch.tensor([1 if torch.equal(a_arr, a_own) else 0], dtype=torch.int32, device=renderer.device)
        if world > 1:
            dist.all_reduce(same_arr, op=dist.ReduceOp.MIN)
        out["gathered_arrays_bitwise"] = bool(same_arr.item())
        arr.close()
        out["replica_bitwise"] = bool(same.item())
        out["replica_built_on_rank0_only"] = bool(rank == 0 or rep.info()["build_ms"] == 0.0)
        rep.close(); own.close()
    return out
