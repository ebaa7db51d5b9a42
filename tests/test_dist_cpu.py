"""Host-side multi-GPU logic on CPU: world_size-2 gloo processes check that the tile partition covers
the frame exactly once, that the single gather (tile sharding) / reduce (sample sharding) reassembles the full
frame on rank 0, and the replicate-instead-of-rebuild protocol including its failure path."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pathtracerpython_b200 import dist as pdist


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, W, H, tile, mode, spp, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.zeros((W * H, 4), dtype=torch.float32)
    if mode == "tile":
        idx = pdist.shard_ray_indices(W, H, rank, world, tile)
        assert len(idx) == pdist.shard_pixel_count(W, H, rank, world, tile)
        # a fake per-pixel "radiance sum" that depends only on the ray index, and the sample count (pixel-slot order)
        vals = torch.tensor(np.stack([idx, 2 * idx, 3 * idx, np.full_like(idx, spp)], 1), dtype=torch.float32)
        full.fill_(-1.0)      # the gather must overwrite every pixel: nothing is summed, nothing needs zeroing

        def scatter(g, shard, dst):      # host stand-in for ptb_scatter_accum
            dst[torch.as_tensor(pdist.shard_ray_indices(W, H, g, world, tile))] = shard
        pdist.gather_frame(vals, W, H, rank, world, tile, full, scatter)
    else:
        s0, s1 = pdist.sample_range(spp, rank, world)
        r = torch.arange(W * H, dtype=torch.float32)
        full[:, 0] = r * (s1 - s0)
        full[:, 3] = float(s1 - s0)
        pdist.reduce_frame(full, 0)
    if rank == 0:
        torch.save(full, out)
    dist.barrier()
    dist.destroy_process_group()


def _run(tmp_path, W, H, tile, mode, spp=5, world=2):
    out = str(tmp_path / "full.pt")
    mp.spawn(_worker, args=(world, _free_port(), W, H, tile, mode, spp, out), nprocs=world, join=True)
    return torch.load(out)


def test_tile_sharding_gather_gloo(tmp_path):
    for (W, H, tile) in [(37, 11, 5), (16, 8, 8)]:      # ragged shards (padded gather) and equal ones
        full = _run(tmp_path, W, H, tile, "tile")
        r = torch.arange(W * H, dtype=torch.float32)
        assert torch.equal(full[:, 0], r) and torch.equal(full[:, 1], 2 * r) and torch.equal(full[:, 3], torch.full((W * H,), 5.0))
    assert pdist.collective_info(1024, 1024, 8, "tile") == {"op": "gather", "bytes": 1024 * 1024 * 16 * 7 // 8,
                                                            "note": "owned shards to rank 0, float4 per pixel"}
    assert pdist.collective_info(64, 64, 1, "tile")["bytes"] == 0


def test_sample_sharding_reduce_gloo(tmp_path):
    W, H = 16, 16
    full = _run(tmp_path, W, H, 0, "sample", spp=7)
    r = torch.arange(W * H, dtype=torch.float32)
    assert torch.equal(full[:, 0], 7 * r) and torch.equal(full[:, 3], torch.full((W * H,), 7.0))


def test_partition_properties():
    for (W, H, tile, G) in [(40, 40, 64, 3), (1024, 1024, 1024, 8), (37, 11, 5, 4), (8, 8, 100, 2)]:
        parts = [pdist.shard_ray_indices(W, H, g, G, tile) for g in range(G)]
        allidx = np.concatenate(parts)
        assert len(allidx) == W * H and len(np.unique(allidx)) == W * H
        # agrees with the C ABI's count
        import ctypes
        from pathtracerpython_b200 import _lib
        for g in range(G):
            px = _lib.Pixels(W, H, g, G, tile, 0, None)
            assert _lib.lib().ptb_pixel_count(ctypes.byref(px)) == len(parts[g])
    assert [pdist.sample_range(10, r, 4) for r in range(4)] == [(0, 2), (2, 5), (5, 7), (7, 10)]


class _FakeRenderer:
    """Stand-in for Renderer in the CPU test of dist.replicate_renderer: 'builds' a blob from the scene."""
    def __init__(self, scene, bvh):
        self.scene, self.imported = scene, bvh is not None
        self.blob = bvh if bvh is not None else (torch.arange(scene["n"], dtype=torch.int64) * 7 % 251).to(torch.uint8)

    def info(self):
        return {"accel": "lbvh" if self.scene["n"] else "brute"}

    def export_bvh(self):
        if self.scene["n"] == 0:
            raise RuntimeError("no LBVH")
        return self.blob


def _replicate_worker(rank, world, port, n, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    made = []

    def make(scene, device, accel, bvh):
        made.append(bvh is not None)
        if scene["n"] < 0:
            raise MemoryError("out of device memory (injected)")
        return _FakeRenderer(scene, bvh)
    try:
        r = pdist.replicate_renderer({"n": n}, "cpu", src=0, make=make)
        torch.save({"imported": r.imported, "made": made, "blob": r.blob}, f"{out}.{rank}")
    except Exception as e:      # noqa: BLE001
        torch.save({"error": type(e).__name__, "made": made}, f"{out}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


def test_replicate_renderer_protocol_gloo(tmp_path):
    """rank 0 builds and broadcasts, rank 1 imports the same bytes; a scene with nothing to export is built everywhere."""
    out = str(tmp_path / "rep")
    mp.spawn(_replicate_worker, args=(2, _free_port(), 1000, out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert r0["made"] == [False] and not r0["imported"]
    assert r1["made"] == [True] and r1["imported"] and torch.equal(r0["blob"], r1["blob"])
    mp.spawn(_replicate_worker, args=(2, _free_port(), 0, out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert r0["made"] == [False] and r1["made"] == [False] and not r1["imported"]
    # the building rank fails: every rank raises (nobody waits in the broadcast), and nobody rebuilds silently
    mp.spawn(_replicate_worker, args=(2, _free_port(), -1, out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert r0["error"] == "MemoryError" and r1["error"] == "RuntimeError" and r1["made"] == []


def _gather_arrays_worker(rank, world, port, n, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(3)
    flat = {"verts": rng.normal(size=(n, 3, 3)), "normals": rng.normal(size=(n, 3)), "tri_obj": rng.integers(0, 7, n).astype(np.int32)}
    v, nr, o = pdist.gather_scene_arrays(flat, "cpu")
    ok = np.array_equal(v.numpy(), flat["verts"]) and np.array_equal(nr.numpy(), flat["normals"]) and np.array_equal(o.numpy(), flat["tri_obj"])
    torch.save({"ok": ok, "shapes": (tuple(v.shape), tuple(nr.shape), tuple(o.shape))}, f"{out}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


def test_scene_arrays_assembled_from_per_rank_slices_gloo(tmp_path):
    """dist.gather_scene_arrays: each rank contributes its rows, one all-gather completes the arrays everywhere; a
    triangle count that does not divide by the rank count goes through the padded tail."""
    out = str(tmp_path / "ga")
    for n in (1001, 64):
        mp.spawn(_gather_arrays_worker, args=(2, _free_port(), n, out), nprocs=2, join=True)
        for r in range(2):
            res = torch.load(f"{out}.{r}")
            assert res["ok"] and res["shapes"] == ((n, 3, 3), (n, 3), (n,))
