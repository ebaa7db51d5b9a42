#!/usr/bin/env python3
"""Run under torchrun with N >= 2 ranks (one per GPU): renders the Cornell box with tile sharding (one NCCL gather)
and with sample-range sharding (one NCCL reduce) through pathtracerpython_b200.dist, and rank 0 checks the result
against its own single-GPU render: tile sharding must be BITWISE identical (every pixel is computed by one rank in
the same order), sample sharding equal up to float32 summation order.  Then a hierarchy built on rank 0 and
broadcast must render bit-identically to a locally built one on every rank.  (bench.py runs the same check before
its timed loop at N > 1 and puts the outcome in its JSON line.)"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pathtracerpython_b200 as ptb  # noqa: E402
from pathtracerpython_b200 import dist as pdist  # noqa: E402
from pathtracerpython_b200 import synth  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    r = ptb.Renderer(ptb.CORNELL_SDL, device=local)
    big = synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 8)
    res = pdist.multi_gpu_check(r, rank, world, lbvh_scene=big)
    # a frame whose width does not divide by the rank count: shards of unequal size through the padded gather
    W, H = 250 + world + 1, 64
    img, full = pdist.render_step(r, 2, 2, 3, W, H, rank, world, "tile", H)
    if rank == 0:
        ref, _ = r.trace_paths(2, 2, seed=3, width=W, height=H)
        assert torch.equal(ref, full), "ragged tile-sharded frame differs from the single-GPU frame"
        assert torch.equal(r.image(ref, W, H), img)
        assert res["tile_bitwise"], "tile-sharded frame differs from the single-GPU frame"
        assert res["sample_max_rel"] < 1e-4 or res["sample_max_abs"] < 1e-4, res
        print(f"multi-GPU check ok on {world} ranks: tile sharding bitwise identical, sample sharding within {res['sample_max_abs']:.2e}")
    assert res["replica_bitwise"] and res["replica_built_on_rank0_only"] and res["gathered_arrays_bitwise"], res
    small = pdist.replicate_renderer(ptb.CORNELL_SDL, local, src=0)      # nothing to export: every rank builds
    assert small.info()["accel"] == "brute"
    if rank == 0:
        print("replicated hierarchy ok")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
