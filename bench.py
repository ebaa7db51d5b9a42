#!/usr/bin/env python3
"""bench.py - headline benchmark of the path-tracing hot path (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--res 1024] [--spp 1024] [--bounces 5] [--scene PATH] [--no-cpu-baseline]

A step = one full frame of the workload (Cornell box, res x res, spp samples per pixel, `bounces`
bounces): camera rays -> wavefront bounces -> accumulation buffer (-> NCCL reduce on N > 1) -> tone map.
`value` = pixel samples per second over the whole job with the scene resident in HBM; `e2e` = the same
through the public call with HOST buffers: flattened scene arrays uploaded, image read back, every step.
N > 1: one process per GPU (torchrun), the frame is tile-sharded (column-interleaved), per-GPU
accumulation buffers are summed with one NCCL reduce to rank 0; fixed total work => "strong" scaling.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"      # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line

F_TEST = 38          # FP32 flops of one ray-triangle test as written in ptb_isect.cuh (FMA = 2); DESIGN.md
RAY_STATE_BYTES = {"closest_hit": 32 + 8, "shadow": 16 + 48 + 8 + 32, "shade": 40 + 32 + 72}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--res", type=int, default=1024)
    ap.add_argument("--width", type=int, default=0, help="frame width (default: --res)")
    ap.add_argument("--height", type=int, default=0, help="frame height (default: --res); config 5 is 3840 x 2160")
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--bounces", type=int, default=5)
    ap.add_argument("--scene", default=None)
    ap.add_argument("--tess", type=int, default=0, help="tessellate every triangle into N x N (configs 4/5: 177 -> 1.0M, 559 -> 10.0M triangles)")
    ap.add_argument("--seed", type=int, default=9)
    ap.add_argument("--shard", default="tile", choices=["tile", "sample"], help="multi-GPU partition (SURVEY.md 8e)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the bounded baseline sample")
    a = ap.parse_args()
    a.width, a.height = a.width or a.res, a.height or a.res
    return a


def workload_name(a):
    t = f"_tess{a.tess}_{32 * a.tess * a.tess}tri" if a.tess else ""
    return f"cornell_box{t}_{a.width}x{a.height}_{a.spp}spp_b{a.bounces}"


# ------------------------------------------------------------------ CPU baseline (oracle port, bounded sample)
def cpu_sample(a, seconds):
    """Times the FP64 C restatement of the reference (oracle/pt_oracle.c, OpenMP over pixels) on a crop
    of the same workload: a centred square of pixels x a few samples, same resolution and bounces."""
    import numpy as np
    from oracle import oracle, scene_port
    import pathtracerpython_b200 as ptb
    if a.tess:
        from pathtracerpython_b200 import synth
        flat = synth.tessellate(ptb.Scene(a.scene or ptb.CORNELL_SDL), a.tess).flatten()
    else:
        flat = scene_port.load_scene(a.scene or ptb.CORNELL_SDL)
    orc = oracle.Oracle(flat, a.width, a.height)
    side = 64 if not a.tess else 4
    lox, loy = (a.width - side) // 2, (a.height - side) // 2
    px = np.array([ix * a.height + iy for ix in range(lox, lox + side) for iy in range(loy, loy + side)], np.int64)
    t = time.perf_counter()
    orc.trace(1, a.bounces, a.seed, pixels=px, details=False)
    dt = time.perf_counter() - t
    spp = int(max(1, min(a.spp, seconds / max(dt, 1e-6))))
    if spp == a.spp and seconds / max(dt, 1e-6) > 1.5 * a.spp:
        # the crop finishes all samples early: widen it (same centre) so the sample is about `seconds` of CPU work
        side = int(min(min(a.width, a.height), side * (seconds / (dt * a.spp)) ** 0.5)) & ~1
        lox, loy = (a.width - side) // 2, (a.height - side) // 2
        px = np.array([ix * a.height + iy for ix in range(lox, lox + side) for iy in range(loy, loy + side)], np.int64)
    t = time.perf_counter()
    out = orc.trace(spp, a.bounces, a.seed, pixels=px, details=False)
    dt = time.perf_counter() - t
    n = len(px) * spp
    return {"value": n / dt / 1e6, "unit": "Msamples/s", "cores": oracle.num_threads(), "kind": "port",
            "sample": f"{side}x{side}-pixel centre crop of the {a.width}x{a.height} frame x {spp} spp, b={a.bounces}, "
                      f"FP64 C restatement of the reference (OpenMP), {dt:.1f} s",
            "mrays_per_s": (out["closest_rays"] + out["shadow_rays"]) / dt / 1e6, "seconds": dt, "samples": n}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step = max(2.0, min(20.0, 120.0 / max(1, a.steps + a.warmup)))
    for _ in range(a.warmup):
        cpu_sample(a, 0.5)
    vals = []
    t0 = time.perf_counter()
    for _ in range(a.steps):
        vals.append(cpu_sample(a, per_step))
    wall = time.perf_counter() - t0
    samples = sum(v["samples"] for v in vals)
    secs = sum(v["seconds"] for v in vals)
    value = samples / secs / 1e6
    cb = dict(vals[-1]); cb["value"] = value
    print(json.dumps({
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * wall / max(1, a.steps), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "note": "CPU reference arm: bounded crop per step, see cpu_baseline.sample"},
        "cpu_baseline": cb,
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(timeout=10)      # an nvidia-smi query still in flight holds a driver lock: keep it out of the e2e leg
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": max((int(r[1]) for r in self.rows if r[1].isdigit()), default=None),
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------ our arm
def run_ours(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    import pathtracerpython_b200 as ptb
    from pathtracerpython_b200 import _lib, dist as pdist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    scene_path = a.scene or ptb.CORNELL_SDL
    scene = ptb.Scene(scene_path)
    if a.tess:
        from pathtracerpython_b200 import synth
        scene = synth.tessellate(scene, a.tess)
    W, H = a.width, a.height
    r = ptb.Renderer(scene, device=local)
    px, _ = r.pixels(W, H, rank, world, H)                    # column-interleaved tiles
    full = torch.zeros((W * H, 4), dtype=torch.float32, device=dev)
    host_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()

    def step(rnd):
        img, _ = pdist.render_step(rnd, a.spp, a.bounces, a.seed, W, H, rank, world, a.shard, H, full)
        return img

    def timed(fn, steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(a.warmup):
        step(r)
    torch.cuda.synchronize()
    r.counters(reset=True)
    _lib.profile_enable(True)
    _lib.profile_read()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = timed(lambda: step(r), a.steps)
    prof = _lib.profile_read()
    _lib.profile_enable(False)
    ctr = r.counters()
    clk = clocks.summary() if rank == 0 else None

    # ---- e2e: host scene arrays (page-locked) in, host image out, every step (scene upload + BVH/records build included)
    scene.flatten(pin=True)

    def e2e_step():
        t0 = time.perf_counter()
        rr = ptb.Renderer(scene, device=local)
        t1 = time.perf_counter()
        img = step(rr)
        if img is not None:
            host_img.copy_(img, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        t2 = time.perf_counter()
        rr.close()
        if os.environ.get("PTB_TIMING"):
            print(f"[bench e2e] create {1e3 * (t1 - t0):.1f} ms, frame + image {1e3 * (t2 - t1):.1f} ms, close {1e3 * (time.perf_counter() - t2):.1f} ms", file=sys.stderr)

    e2e_step()
    ms_e2e = timed(e2e_step, a.steps)
    flat = r.flat
    h2d = int(sum(np.asarray(flat[k]).nbytes for k in ("verts", "normals", "tri_obj", "mat", "light_areas")))

    rays = torch.tensor([ctr["closest_rays"] + ctr["shadow_rays"], ctr["closest_rays"], ctr["shadow_rays"], ctr["shaded"]],
                        dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(rays)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    samples = W * H * a.spp * a.steps
    sec = ms / 1e3
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    fp32_peak = _lib.fp32_peak_tflops() if hasattr(_lib, "fp32_peak_tflops") else None
    # dominant kernel of this rank
    top = max(("closest_hit", "shadow", "shade", "bounce_fused"), key=lambda k: prof[k][0])
    n_tri, n_tri_obj = r.n_tri, int(flat["n_tri_obj"])
    t_top = prof[top][0] / 1e3
    info = r.info()
    kernel_ms = {k: round(v[0], 3) for k, v in prof.items() if v[1]}
    if info["accel"] == "lbvh":
        # instrumented, untimed pass of the same kernels on the same LBVH: nodes fetched / triangles tested per ray
        r.counters(reset=True)
        r.trace_paths(1, a.bounces, seed=a.seed, width=W, height=H, shard=rank if a.shard == "tile" else 0,
                      n_shards=world if a.shard == "tile" else 1, tile=H, flags=_lib.FLAG_COUNT_NODES)
        ci = r.counters(reset=True)
        n_ch, t_ch = ci["nodes_visited"] / max(1, ci["closest_rays"]), ci["tris_tested"] / max(1, ci["closest_rays"])
        n_sh, t_sh = ci["nodes_visited_shadow"] / max(1, ci["shadow_rays"]), ci["tris_tested_shadow"] / max(1, ci["shadow_rays"])
        if top == "shadow":
            bytes_ = ctr["shadow_rays"] * (64 * n_sh + 64 * t_sh) + ctr["shaded"] * RAY_STATE_BYTES["shadow"]
            flops = ctr["shadow_rays"] * (36 * n_sh + F_TEST * t_sh)
        else:
            top = "closest_hit"
            t_top = prof[top][0] / 1e3
            bytes_ = ctr["closest_rays"] * (RAY_STATE_BYTES["closest_hit"] + 64 * n_ch + 64 * t_ch)
            flops = ctr["closest_rays"] * (36 * n_ch + F_TEST * t_ch)
        roof = {"kernel": top, "bound": "hbm", "achieved": bytes_ / t_top / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": bytes_ / t_top / 1e9 / hbm_peak,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650",
                "traffic": None,
                "per_ray": {"nodes_closest": n_ch, "tris_closest": t_ch, "nodes_shadow": n_sh, "tris_shadow": t_sh},
                "fp32": {"achieved": flops / t_top / 1e12, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": flops / t_top / 1e12 / fp32_peak if fp32_peak else None},
                "launches": prof[top][1], "avg_launch_ms": prof[top][0] / max(1, prof[top][1]), "kernel_ms": kernel_ms,
                "note": "algorithmic bytes = 64 B per node fetched + 64 B per triangle record tested + ray/hit state (DESIGN.md)"}
    else:
        if top == "bounce_fused":
            # one kernel per bounce: 3 shadow rays over the occlusion set + the next closest hit over all triangles
            # (per-triangle formulation) + shading; the per-pixel primary rays (< 1 % of closest rays)
            # are traced by the separate tracked kernel and are left in the count
            flops = (ctr["shadow_rays"] * n_tri_obj + ctr["closest_rays"] * n_tri) * F_TEST + ctr["shaded"] * 400
            bytes_ = ctr["shaded"] * (40 + 32) + ctr["closest_rays"] * 40
        elif top == "closest_hit":
            flops = ctr["closest_rays"] * n_tri * F_TEST
            bytes_ = ctr["closest_rays"] * RAY_STATE_BYTES["closest_hit"]
        elif top == "shadow":
            flops = ctr["shadow_rays"] * n_tri_obj * F_TEST     # the per-triangle formulation's work
            bytes_ = ctr["shaded"] * RAY_STATE_BYTES["shadow"]
        else:
            flops = ctr["shaded"] * 400
            bytes_ = ctr["closest_rays"] * RAY_STATE_BYTES["shade"]
        roof = {"kernel": top, "bound": "fp32", "achieved": flops / t_top / 1e12 if t_top else None,
                "peak": fp32_peak, "unit": "TFLOP/s",
                "frac": (flops / t_top / 1e12 / fp32_peak) if (t_top and fp32_peak) else None,
                "peak_source": "FP32 FMA microbenchmark run by this bench (ptb_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
                "traffic": None,
                "hbm": {"achieved": bytes_ / t_top / 1e9 if t_top else None, "peak": hbm_peak, "unit": "GB/s",
                        "frac": bytes_ / t_top / 1e9 / hbm_peak if t_top else None,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650"},
                "launches": prof[top][1], "avg_launch_ms": prof[top][0] / max(1, prof[top][1]), "kernel_ms": kernel_ms,
                "note": "scene lives in shared memory: the kernel is FP32-issue bound, not HBM bound (DESIGN.md)"}
    # DRAM bytes per launch of the dominant kernel: measured per-unit figure from the committed ncu pass
    # (profiles/r01r_traffic.json; ncu cannot run inside a timed bench) x the units one launch processed here
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r01r_traffic.json")))
        key, units = {"bounce_fused": ("bounce_fused", ctr["shaded"]), "shadow": ("shadow_lbvh", ctr["shadow_rays"]),
                      "closest_hit": ("closest_lbvh", ctr["closest_rays"])}.get(roof["kernel"], (None, 0))
        if key and (key == "bounce_fused") == (info["accel"] == "brute") and roof["launches"]:
            roof["traffic"] = tj[key]["dram_bytes_per_unit"] * units / roof["launches"]
            roof["traffic_source"] = f"profiles/r01r_traffic.json: {tj[key]['dram_bytes_per_unit']} B per {tj[key]['unit']} (ncu dram__bytes_read+write) x units per launch of this run"
            roof["algorithmic_bytes_per_launch"] = bytes_ / roof["launches"]
    except Exception:
        pass
    line = {
        "metric": "Msamples/s", "value": samples / sec / 1e6, "unit": "Msamples/s", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a), "scene": os.path.basename(scene_path), "triangles": n_tri,
                   "accel": info["accel"], "bvh_build_ms": info["build_ms"],
                   "width": W, "height": H, "spp": a.spp, "bounces": a.bounces, "sharding": (f"column-interleaved tiles x{world}" if a.shard == "tile" else f"sample ranges x{world}"),
                   "l2_policy": "inputs larger than L2 (wavefront state of one batch is GBs)"},
        "mrays_per_s": float(rays[0]) / sec / 1e6, "closest_mrays_per_s": float(rays[1]) / sec / 1e6,
        "shadow_mrays_per_s": float(rays[2]) / sec / 1e6, "rays_per_sample": float(rays[0]) / samples,
        "ms_per_frame": ms / a.steps,
        "e2e": {"value": samples / (ms_e2e / 1e3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": W * H * 3, "ms_per_step": ms_e2e / a.steps},
        "gpu_launches": ctr["launches"], "roofline": roof, "clocks": clk,
    }
    if not a.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_sample(a, a.cpu_seconds)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
