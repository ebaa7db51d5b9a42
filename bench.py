#!/usr/bin/env python3
"""bench.py - headline benchmark of the path-tracing hot path (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--res 1024] [--spp 1024] [--bounces 5] [--scene PATH] [--tess N] [--warp] [--no-secondary]

A step = one full frame of the workload (Cornell box, res x res, spp samples per pixel, `bounces`
bounces): camera rays -> wavefront bounces -> accumulation buffer (-> one NCCL collective on N > 1) -> tone map.
`value` = pixel samples per second over the whole job with the scene resident in HBM; `e2e` = the same
through the public call with HOST buffers: flattened scene arrays uploaded, image read back, every step.
N > 1: one process per GPU (torchrun); the frame is tile-sharded (column-interleaved) and the owned shards are
gathered on rank 0 (Cornell), or sample-range-sharded with one sum-reduce (config 5); fixed total work => "strong".

The ONE JSON line carries the headline workload (BASELINE.json configs[2]: Cornell 1024^2 x 1024 spp) and, under
`secondary`, the other named configurations measured in the same run with their own small step counts
(configs[1] Cornell 512^2 x 64 spp; configs[3] the 1 M-triangle tessellation; a WARPED 1 M-triangle scene where the
supporting-plane culling cannot apply; configs[4]'s 10 M-triangle scene per GPU and, at N > 1, sample-sharded).
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
FP32_NOMINAL_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12      # 148 SMs x 128 FMA lanes x 2 flop x 1965 MHz = 74.4
RAY_STATE_BYTES = {"closest_hit": 32 + 8, "shadow": 16 + 48 + 8 + 32, "shade": 40 + 32 + 72}
TRAFFIC_FILES = ["r02_traffic.json", "r01r_traffic.json"]       # measured DRAM bytes per unit (ncu), newest first


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
    ap.add_argument("--warp", action="store_true", help="warp the (tessellated) scene so that no two triangles are coplanar")
    ap.add_argument("--seed", type=int, default=9)
    ap.add_argument("--shard", default="tile", choices=["tile", "sample"], help="multi-GPU partition (SURVEY.md 8e)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="headline workload only")
    ap.add_argument("--secondary-steps", type=int, default=2)
    ap.add_argument("--no-e2e-variants", action="store_true", help="N > 1, LBVH scenes: time only the per-rank-upload end-to-end step")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the bounded baseline sample")
    ap.add_argument("--config1", action="store_true", help="BASELINE.json configs[0]: the shipped 40x40 scene at -r 10 -b 2, "
                    "GPU next to the UNMODIFIED Python reference (oracle/_ref) on this host")
    a = ap.parse_args()
    a.width, a.height = a.width or a.res, a.height or a.res
    return a


class Spec:
    """One workload."""
    def __init__(self, width, height, spp, bounces, tess=0, warp=False, shard="tile", seed=9, scene=None, label=None):
        self.width, self.height, self.spp, self.bounces = width, height, spp, bounces
        self.tess, self.warp, self.shard, self.seed, self.scene, self.label = tess, warp, shard, seed, scene, label

    @property
    def name(self):
        t = f"_tess{self.tess}_{32 * self.tess * self.tess}tri" if self.tess else ""
        w = "_warped" if self.warp else ""
        return f"cornell_box{t}{w}_{self.width}x{self.height}_{self.spp}spp_b{self.bounces}"

    def n_tri(self):
        return 32 * (self.tess * self.tess if self.tess else 1)

    def config(self, world):
        """Identical for both arms (ours / reference): only what defines the workload."""
        return {"workload": self.name, "scene": os.path.basename(self.scene) if self.scene else "cornellroom.sdl",
                "triangles": self.n_tri() if not self.scene else None, "width": self.width, "height": self.height,
                "spp": self.spp, "bounces": self.bounces, "seed": self.seed,
                "sharding": (f"column-interleaved tiles x{world}" if self.shard == "tile" else f"sample ranges x{world}"),
                "l2_policy": "inputs larger than L2 (wavefront state of one batch is GBs)"}

    def build_scene(self):
        import pathtracerpython_b200 as ptb
        scene = ptb.Scene(self.scene or ptb.CORNELL_SDL)
        if self.tess:
            from pathtracerpython_b200 import synth
            scene = synth.tessellate(scene, self.tess)
        if self.warp:
            from pathtracerpython_b200 import synth
            scene = synth.warp(scene)
        return scene


def spec_of_args(a):
    return Spec(a.width, a.height, a.spp, a.bounces, a.tess, a.warp, a.shard, a.seed, a.scene)


# ------------------------------------------------------------------ CPU baseline (oracle port, bounded sample)
def cpu_sample(spec, seconds, threads=0):
    """Times the FP64 C restatement of the reference (oracle/pt_oracle.c, OpenMP) on a crop of the same
    workload: a centred square of pixels x a few samples, same resolution and bounces, on ALL the host cores this
    process may use (torchrun's OMP_NUM_THREADS=1 is overridden: oracle.set_threads)."""
    import numpy as np
    from oracle import oracle, scene_port
    import pathtracerpython_b200 as ptb
    cores = oracle.set_threads(threads)
    if spec.tess or spec.warp:
        flat = spec.build_scene().flatten()
    else:
        flat = scene_port.load_scene(spec.scene or ptb.CORNELL_SDL)
    orc = oracle.Oracle(flat, spec.width, spec.height)
    W, H = spec.width, spec.height

    def crop(side):
        lox, loy = (W - side) // 2, (H - side) // 2
        return np.array([ix * H + iy for ix in range(lox, lox + side) for iy in range(loy, loy + side)], np.int64)
    side = 64 if not spec.tess else 4
    px = crop(side)
    t = time.perf_counter()
    orc.trace(1, spec.bounces, spec.seed, pixels=px, details=False)
    dt = time.perf_counter() - t
    spp = int(max(1, min(spec.spp, seconds / max(dt, 1e-6))))
    if spp == spec.spp and seconds / max(dt, 1e-6) > 1.5 * spec.spp:
        # the crop finishes all samples early: widen it (same centre) so the sample is about `seconds` of CPU work
        side = int(min(min(W, H), side * (seconds / (dt * spec.spp)) ** 0.5)) & ~1
        px = crop(side)
    t = time.perf_counter()
    out = orc.trace(spp, spec.bounces, spec.seed, pixels=px, details=False)
    dt = time.perf_counter() - t
    n = len(px) * spp
    return {"value": n / dt / 1e6, "unit": "Msamples/s", "cores": cores, "host_cores": oracle.host_cores(), "kind": "port",
            "sample": f"{side}x{side}-pixel centre crop of the {W}x{H} frame x {spp} spp, b={spec.bounces}, "
                      f"FP64 C restatement of the reference (OpenMP, {cores} threads), {dt:.1f} s",
            "mrays_per_s": (out["closest_rays"] + out["shadow_rays"]) / dt / 1e6, "seconds": dt, "samples": n}


def python_reference_sample(n_rays=1, n_bounces=2):
    """The UNMODIFIED Python reference (oracle/_ref, copied by oracle/make_ref.py) on the shipped 40x40 scene through
    its own Pool(cpu_count()) (main.py:197, 208): SURVEY.md 8(d) CPU baseline (i)."""
    from oracle import make_ref
    if not make_ref.available():
        return {"unavailable": "oracle/_ref is empty (run oracle/make_ref.py where /root/reference is mounted)"}
    try:
        r = make_ref.run_main(n_rays, n_bounces)
    except Exception as e:      # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    return {"value": r["samples_per_s"] / 1e6, "unit": "Msamples/s", "samples_per_s": r["samples_per_s"], "cores": r["cores"],
            "kind": "reference", "seconds": r["seconds"], "samples": r["samples"],
            "sample": f"unmodified reference, {r['command']} (40x40 as shipped), multiprocessing.Pool({r['cores']}), {r['seconds']:.1f} s"}


def parity_sample(spec_small, renderer):
    """Checker leg (never timed): per-path records of the product path against the oracle on a small frame of the same
    scene -> measured divergent fraction (path-bounces whose discrete history differs) in ppm."""
    import numpy as np
    from oracle import oracle
    flat = renderer.flat
    W, H, spp, nb, seed = spec_small
    ref = oracle.Oracle(flat, W, H).trace(spp, nb, seed)
    _, rec = renderer.trace_paths(spp, nb, seed=seed, width=W, height=H, records=True)
    rec = {k: v.cpu().numpy() for k, v in rec.items()}
    same = (rec["tri"] == ref["tri"]) & (rec["lobe"] == ref["lobe"]) & (rec["occ"] == ref["occ"]) & (rec["leak"] == ref["leak"])
    ok = np.logical_and.accumulate(same, axis=2)
    live = ref["tri"] > -2
    n_live = int(live.sum())
    bad_hist = int((~ok & live).sum())
    tol = np.abs(rec["contrib"] - ref["contrib"]) <= 1e-4 + 1e-3 * np.abs(ref["contrib"])
    return {"frame": f"{W}x{H}x{spp}spp_b{nb}", "path_bounces": n_live, "divergent_path_bounces": bad_hist,
            "divergent_ppm": 1e6 * bad_hist / max(1, n_live),
            "values_outside_tolerance_on_matching_histories": int((~tol)[ok].sum()),
            "tolerance": "|d| <= 1e-4 + 1e-3 |ref| on contrib (FP32 product vs FP64 oracle, identical Philox stream)"}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    spec = spec_of_args(a)
    world = int(os.environ.get("WORLD_SIZE", str(a.gpus)))
    per_step = max(2.0, min(20.0, 120.0 / max(1, a.steps + a.warmup)))
    for _ in range(a.warmup):
        cpu_sample(spec, 0.5)
    vals = []
    t0 = time.perf_counter()
    for _ in range(a.steps):
        vals.append(cpu_sample(spec, per_step))
    wall = time.perf_counter() - t0
    samples = sum(v["samples"] for v in vals)
    secs = sum(v["seconds"] for v in vals)
    value = samples / secs / 1e6
    cb = dict(vals[-1]); cb["value"] = value
    cb["all_cores"] = bool(cb["cores"] == cb["host_cores"])      # torchrun's OMP_NUM_THREADS=1 no longer applies
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * wall / max(1, a.steps), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": spec.config(world),
        "note": "CPU reference arm: the reference's algorithm (FP64 C restatement, all host cores) on a bounded crop of the "
                "workload per step, see cpu_baseline.sample; rank 0 only",
        "cpu_baseline": cb,
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of one GPU, sampled every 0.2 s DURING the timed region.  In-process NVML queries
    (nvidia_ml_py: the library nvidia-smi itself is a client of) - spawning nvidia-smi five times a second was measured to
    stall the GPU's queue for milliseconds at a time, which a 4 ms frame cannot absorb; nvidia-smi stays as the fallback."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.nvml, self.handle = None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml, self.handle = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self._nvml_row()
        except Exception:
            self.nvml = None

    def _nvml_row(self):
        N, h = self.nvml, self.handle
        sm, mx = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM), N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
        try:
            r = N.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            r = N.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        bits = (N.nvmlClocksThrottleReasonHwSlowdown, N.nvmlClocksThrottleReasonHwThermalSlowdown,
                N.nvmlClocksThrottleReasonSwThermalSlowdown, N.nvmlClocksThrottleReasonSwPowerCap)
        return [str(sm), str(mx)] + ["Active" if (r & b) else "Not Active" for b in bits]

    def run(self):
        while not self.stop_flag:
            try:
                if self.nvml is not None:
                    self.rows.append(self._nvml_row())
                else:
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
            self.join(timeout=10)      # a query still in flight holds a driver lock: keep it out of the e2e leg
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": max((int(r[1]) for r in self.rows if r[1].isdigit()), default=None),
                "reasons": reasons, "samples": len(self.rows), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


# ------------------------------------------------------------------ our arm
class Ctx:
    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self._peaks = None

    def peaks(self):
        """Roofline denominators: MEASURED_PEAKS.json (driver-written) + the library's own microbenchmarks for what it
        does not hold (FP32 FMA issue, 64 B gathers out of L1 / L2 / HBM)."""
        if self._peaks is None:
            from pathtracerpython_b200 import _lib
            p = {}
            try:
                p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            except Exception:
                pass
            self._peaks = {"hbm_gbs": p.get("hbm_gbs", 6650.0),
                           "hbm_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in p else "fallback 6650",
                           "fp32_tflops": _lib.fp32_peak_tflops(),
                           "gather_l1_gfetch": _lib.gather_peak_gfetch(64 << 10),
                           "gather_l2_gfetch": _lib.gather_peak_gfetch(64 << 20),
                           "gather_hbm_gfetch": _lib.gather_peak_gfetch(1536 << 20)}
        return self._peaks

    def timed(self, fn, steps):
        torch, dist = self.torch, self.dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())


def traffic_table():
    for f in TRAFFIC_FILES:
        try:
            return json.load(open(os.path.join(ROOT, "profiles", f))), f
        except Exception:
            continue
    return None, None


def measure(ctx, spec, steps, warmup, e2e_replicated=False, min_region_ms=0.0):
    """One workload on this rank's share: resident-scene value, end-to-end value, ray rates, roofline of the dominant
    kernel, clocks.  Every rank calls it; the dict is meaningful on rank 0."""
    import numpy as np
    import pathtracerpython_b200 as ptb
    from pathtracerpython_b200 import _lib, dist as pdist
    torch, dist, world, rank, local, dev = ctx.torch, ctx.dist, ctx.world, ctx.rank, ctx.local, ctx.dev
    scene = spec.build_scene()
    W, H = spec.width, spec.height
    r = ptb.Renderer(scene, device=local)
    full = torch.zeros((W * H, 4), dtype=torch.float32, device=dev) if (spec.shard == "sample" or rank == 0) else None
    host_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()

    def step(rnd):
        img, _ = pdist.render_step(rnd, spec.spp, spec.bounces, spec.seed, W, H, rank, world, spec.shard, H, full)
        return img

    for _ in range(warmup):
        step(r)
    torch.cuda.synchronize()
    if min_region_ms > 0:
        # secondary workloads with millisecond frames: time enough frames for the region to outlast one scheduling hiccup
        # or one nvidia-smi query of the clock sampler (a 512^2 x 64 spp frame is 3.8 ms; the headline keeps EXACTLY `steps`)
        est = ctx.timed(lambda: step(r), 1)
        steps = int(max(steps, min(2000, -(-min_region_ms // max(est, 1e-3)))))
    r.counters(reset=True)
    _lib.profile_enable(True)
    _lib.profile_read()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = ctx.timed(lambda: step(r), steps)
    prof = _lib.profile_read()
    _lib.profile_enable(False)
    ctr = r.counters()
    clk = clocks.summary() if rank == 0 else None

    # ---- e2e: host scene arrays (page-locked) in, host image out, every step (scene upload + BVH/records build included)
    scene.flatten(pin=True)

    def e2e_step(mode="own"):
        t0 = time.perf_counter()
        if mode == "replicate":
            rr = pdist.replicate_renderer(scene, local, src=0)
        elif mode == "gather":
            rr = ptb.Renderer(scene, device=local, device_arrays=pdist.gather_scene_arrays(scene, dev, rank, world))
        else:
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
    ms_e2e = ctx.timed(e2e_step, steps)
    ms_e2e_rep = ms_e2e_gat = None
    if e2e_replicated and world > 1 and r.info()["accel"] == "lbvh":
        e2e_step("replicate")
        ms_e2e_rep = ctx.timed(lambda: e2e_step("replicate"), steps)
        e2e_step("gather")
        ms_e2e_gat = ctx.timed(lambda: e2e_step("gather"), steps)
    flat = r.flat
    h2d = int(sum(np.asarray(flat[k]).nbytes for k in ("verts", "normals", "tri_obj", "mat", "light_areas")))

    rays = torch.tensor([ctr["closest_rays"] + ctr["shadow_rays"], ctr["closest_rays"], ctr["shadow_rays"], ctr["shaded"]],
                        dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(rays)
    info = r.info()
    entry = None
    if rank == 0:
        samples = W * H * spec.spp * steps
        sec = ms / 1e3
        roof = roofline(ctx, spec, r, prof, ctr, info)
        e2e = {"value": samples / (ms_e2e / 1e3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": W * H * 3, "ms_per_step": ms_e2e / steps,
               "scene": "every rank uploads the host scene arrays and builds its own hierarchies"}
        if ms_e2e_rep is not None:
            e2e["variants"] = {
                "per_rank_upload": {"value": e2e["value"], "ms_per_step": e2e["ms_per_step"], "h2d_bytes_per_step": h2d, "scene": e2e["scene"]},
                "replicated_hierarchy": {"value": samples / (ms_e2e_rep / 1e3) / 1e6, "ms_per_step": ms_e2e_rep / steps, "h2d_bytes_per_step": h2d,
                                         "scene": "rank 0 builds, both hierarchies travel in one NCCL broadcast (dist.replicate_renderer); every rank still uploads the FP64 arrays"},
                "gathered_arrays": {"value": samples / (ms_e2e_gat / 1e3) / 1e6, "ms_per_step": ms_e2e_gat / steps, "h2d_bytes_per_step": h2d // world + 4096,
                                    "scene": "every rank uploads 1/N of the scene arrays, one NCCL all-gather completes them over NVLink, every rank builds (dist.gather_scene_arrays)"}}
            best = max(e2e["variants"], key=lambda k: e2e["variants"][k]["value"])
            e2e.update(value=e2e["variants"][best]["value"], ms_per_step=e2e["variants"][best]["ms_per_step"],
                       h2d_bytes_per_step=e2e["variants"][best]["h2d_bytes_per_step"], scene=e2e["variants"][best]["scene"], variant=best)
        entry = {
            "workload": spec.name, "label": spec.label, "value": samples / sec / 1e6, "unit": "Msamples/s",
            "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
            "scaling": "strong", "config": spec.config(world), "accel": info["accel"], "plane_culling": info["plane_culling"],
            "bvh_build_ms": info["build_ms"],
            "mrays_per_s": float(rays[0]) / sec / 1e6, "closest_mrays_per_s": float(rays[1]) / sec / 1e6,
            "shadow_mrays_per_s": float(rays[2]) / sec / 1e6, "rays_per_sample": float(rays[0]) / samples,
            "e2e": e2e, "gpu_launches": ctr["launches"], "roofline": roof, "clocks": clk,
            "collective": pdist.collective_info(W, H, world, spec.shard)}
    return entry, r, scene


def roofline(ctx, spec, r, prof, ctr, info):
    """Roofline of the dominant kernel class of this rank's timed region (per-kernel CUDA-event times, ptb_profile_*)."""
    from pathtracerpython_b200 import _lib
    pk = ctx.peaks()
    flat = r.flat
    n_tri, n_tri_obj = r.n_tri, int(flat["n_tri_obj"])
    top = max(("closest_hit", "shadow", "shade", "bounce_fused"), key=lambda k: prof[k][0])
    t_top = prof[top][0] / 1e3
    kernel_ms = {k: round(v[0], 3) for k, v in prof.items() if v[1]}
    tj, tj_name = traffic_table()
    if info["accel"] == "lbvh":
        # instrumented, untimed pass of the same kernels on the same LBVH: nodes fetched / triangles tested per ray
        r.counters(reset=True)
        r.trace_paths(1, spec.bounces, seed=spec.seed, width=spec.width, height=spec.height,
                      shard=ctx.rank if spec.shard == "tile" else 0, n_shards=ctx.world if spec.shard == "tile" else 1,
                      tile=spec.height, flags=_lib.FLAG_COUNT_NODES)
        ci = r.counters(reset=True)
        n_ch, t_ch = ci["nodes_visited"] / max(1, ci["closest_rays"]), ci["tris_tested"] / max(1, ci["closest_rays"])
        n_sh, t_sh = ci["nodes_visited_shadow"] / max(1, ci["shadow_rays"]), ci["tris_tested_shadow"] / max(1, ci["shadow_rays"])
        per_kernel = {}
        for kname, nrays, nn, nt, state in (("closest_hit", ctr["closest_rays"], n_ch, t_ch, RAY_STATE_BYTES["closest_hit"]),
                                            ("shadow", ctr["shadow_rays"], n_sh, t_sh, RAY_STATE_BYTES["shadow"] / 3.0)):
            t_k = prof[kname][0] / 1e3
            if not t_k:
                continue
            fetches = nrays * (nn + nt)
            d = {"fetches_per_ray": nn + nt, "gfetch_per_s": fetches / t_k / 1e9, "ms": prof[kname][0],
                 "frac_of_l1_gather_peak": fetches / t_k / 1e9 / pk["gather_l1_gfetch"],
                 "frac_of_l2_gather_peak": fetches / t_k / 1e9 / pk["gather_l2_gfetch"],
                 "algorithmic_bytes": nrays * (64 * (nn + nt) + state)}
            key = {"closest_hit": "closest_lbvh", "shadow": "shadow_lbvh"}[kname]
            if tj and key in tj and info["plane_culling"]:
                dram = tj[key]["dram_bytes_per_unit"] * nrays
                d["dram_bytes"] = dram
                d["dram_gbs"] = dram / t_k / 1e9
                d["dram_frac_of_hbm_peak"] = dram / t_k / 1e9 / pk["hbm_gbs"]
            per_kernel[kname] = d
        if top not in per_kernel:
            top = "shadow" if "shadow" in per_kernel else "closest_hit"
        tk = per_kernel[top]
        launches = prof[top][1]
        roof = {"kernel": top, "bound": "gather",
                "achieved": tk["gfetch_per_s"], "peak": pk["gather_l1_gfetch"], "unit": "Gfetch/s (64 B records)",
                "frac": tk["frac_of_l1_gather_peak"],
                "peak_source": "ptb_gather_peak: dependent random 64 B fetches (2 x LDG.E.256 per lane) at the traversal kernels' "
                               "launch shape; 64 KB buffer = every fetch an L1 hit, the ceiling of any node walk; "
                               "MEASURED_PEAKS.json has no gather figure",
                "peaks": {"l1_64KB_gfetch": pk["gather_l1_gfetch"], "l2_64MB_gfetch": pk["gather_l2_gfetch"],
                          "hbm_1.5GB_gfetch": pk["gather_hbm_gfetch"], "hbm_gbs": pk["hbm_gbs"]},
                "traffic": (tk.get("dram_bytes") / launches) if (tk.get("dram_bytes") and launches) else None,
                "traffic_source": f"profiles/{tj_name}: measured DRAM bytes per ray (ncu dram__bytes_read+write) x rays per launch of this run" if tk.get("dram_bytes") else None,
                "algorithmic_bytes_per_launch": tk["algorithmic_bytes"] / max(1, launches),
                "hbm": {"achieved": tk.get("dram_gbs"), "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": tk.get("dram_frac_of_hbm_peak"),
                        "peak_source": pk["hbm_source"], "note": "MEASURED DRAM traffic, not algorithmic bytes: the trees are served from L1/L2"},
                "per_ray": {"nodes_closest": n_ch, "tris_closest": t_ch, "nodes_shadow": n_sh, "tris_shadow": t_sh},
                "executed": (tj or {}).get("lbvh_executed"),
                "kernels": per_kernel, "launches": launches, "avg_launch_ms": prof[top][0] / max(1, launches), "kernel_ms": kernel_ms,
                "note": "bound is 'gather', outside the contract's {hbm, tensor}: these kernels fetch 64 B nodes / records that live "
                        "in L1 / L2, DRAM sees a few per cent of the fetched bytes (hbm.frac); frac = fetch rate / L1-resident gather peak"}
        return roof
    if top == "bounce_fused":
        # one kernel per bounce: 3 shadow rays over the occlusion set + the next closest hit over all triangles
        # (per-triangle formulation) + shading; the per-pixel primary rays (< 1 % of closest rays)
        # are traced by the separate tracked kernel and are left in the count
        flops = (ctr["shadow_rays"] * n_tri_obj + ctr["closest_rays"] * n_tri) * F_TEST + ctr["shaded"] * 400
        bytes_ = ctr["shaded"] * (40 + 32) + ctr["closest_rays"] * 40
        unit_key, units = "bounce_fused", ctr["shaded"]
    elif top == "closest_hit":
        flops = ctr["closest_rays"] * n_tri * F_TEST
        bytes_ = ctr["closest_rays"] * RAY_STATE_BYTES["closest_hit"]
        unit_key, units = None, 0
    elif top == "shadow":
        flops = ctr["shadow_rays"] * n_tri_obj * F_TEST     # the per-triangle formulation's work
        bytes_ = ctr["shaded"] * RAY_STATE_BYTES["shadow"]
        unit_key, units = None, 0
    else:
        flops = ctr["shaded"] * 400
        bytes_ = ctr["closest_rays"] * RAY_STATE_BYTES["shade"]
        unit_key, units = None, 0
    fp32_peak = pk["fp32_tflops"]
    launches = prof[top][1]
    ach = flops / t_top / 1e12 if t_top else None
    roof = {"kernel": top, "bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s",
            "frac": (ach / fp32_peak) if (ach and fp32_peak) else None,
            "algorithmic_frac": (ach / fp32_peak) if (ach and fp32_peak) else None,
            "peak_source": "FP32 FMA microbenchmark run by this bench (ptb_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
            "peak_nominal": FP32_NOMINAL_TFLOPS, "frac_of_nominal": (ach / FP32_NOMINAL_TFLOPS) if ach else None,
            "executed": (tj or {}).get("bounce_fused_executed") if top == "bounce_fused" else None,
            "traffic": None,
            "hbm": {"achieved": bytes_ / t_top / 1e9 if t_top else None, "peak": pk["hbm_gbs"], "unit": "GB/s",
                    "frac": bytes_ / t_top / 1e9 / pk["hbm_gbs"] if t_top else None, "peak_source": pk["hbm_source"]},
            "launches": launches, "avg_launch_ms": prof[top][0] / max(1, launches), "kernel_ms": kernel_ms,
            "note": "scene lives in shared memory: the kernel is FP32-issue bound, not HBM bound (DESIGN.md).  frac / algorithmic_frac "
                    "= ALGORITHMIC flops of the one-test-per-triangle formulation / measured FMA peak, not pipe utilisation: the "
                    "kernel executes fewer instructions than that (plane sharing, parallelogram test); `executed` holds the ncu "
                    "issue-slot and pipe figures of the committed capture"}
    if tj and unit_key and unit_key in tj and launches:
        roof["traffic"] = tj[unit_key]["dram_bytes_per_unit"] * units / launches
        roof["traffic_source"] = f"profiles/{tj_name}: {tj[unit_key]['dram_bytes_per_unit']} B per {tj[unit_key]['unit']} (ncu dram__bytes_read+write) x units per launch of this run"
        roof["algorithmic_bytes_per_launch"] = bytes_ / launches
    return roof


def run_ours(a):
    ctx = Ctx()
    dist, world, rank = ctx.dist, ctx.world, ctx.rank
    from pathtracerpython_b200 import dist as pdist
    import pathtracerpython_b200 as ptb
    main_spec = spec_of_args(a)
    mgc = None
    if world > 1:
        # correctness of the N-GPU frame, before anything is timed: the sharded frame against rank 0's own 1-GPU frame
        from pathtracerpython_b200 import synth
        r0 = ptb.Renderer(ptb.CORNELL_SDL, device=ctx.local)
        mgc = pdist.multi_gpu_check(r0, rank, world, lbvh_scene=synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 8))
        r0.close()
    entry, r, _ = measure(ctx, main_spec, a.steps, a.warmup, e2e_replicated=bool(a.tess) and not a.no_e2e_variants)
    secondary = []
    if not a.no_secondary and not (a.tess or a.warp or a.scene) and (a.width, a.height, a.spp) == (1024, 1024, 1024):
        ks, kw = max(1, min(a.secondary_steps, a.steps)), 3
        if world == 1:
            specs = [Spec(512, 512, 64, 3, label="C2: BASELINE.json configs[1]"),
                     Spec(1024, 1024, 256, 5, tess=177, label="C4: BASELINE.json configs[3], 1 002 528 triangles in 32 planes (best case for plane culling)"),
                     Spec(1024, 1024, 64, 5, tess=177, warp=True, label="C4w: the same mesh warped, 1 002 528 distinct planes (no plane culling)"),
                     Spec(3840, 2160, 8, 5, tess=559, shard="sample", label="C5 per GPU: BASELINE.json configs[4]'s scene and frame at 8 spp")]
        else:
            specs = [Spec(3840, 2160, 64, 5, tess=559, shard="sample", label=f"C5 strong: 64 spp in total, sample ranges over {world} GPUs"),
                     Spec(3840, 2160, 8 * world, 5, tess=559, shard="sample", label=f"C5 weak: 8 spp per GPU ({8 * world} in total)")]
        for sp in specs:
            e2, r2, _ = measure(ctx, sp, ks, kw, e2e_replicated=True, min_region_ms=300.0)
            if e2 is not None:
                if "weak" in (sp.label or ""):
                    e2["scaling"] = "weak"
                if world == 1 and not a.no_cpu_baseline:
                    try:
                        e2["cpu_baseline"] = cpu_sample(sp, 4.0)
                    except Exception as ex:      # noqa: BLE001
                        e2["cpu_baseline"] = {"unavailable": f"{type(ex).__name__}: {ex}"[:200]}
                secondary.append(e2)
            r2.close()
            del r2
            ptb.release_workspace()
    if rank == 0:
        line = {
            "metric": "Msamples/s", "value": entry["value"], "unit": "Msamples/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": entry["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": entry["config"],
            "accel": entry["accel"], "plane_culling": entry["plane_culling"], "bvh_build_ms": entry["bvh_build_ms"],
            "mrays_per_s": entry["mrays_per_s"], "closest_mrays_per_s": entry["closest_mrays_per_s"],
            "shadow_mrays_per_s": entry["shadow_mrays_per_s"], "rays_per_sample": entry["rays_per_sample"],
            "ms_per_frame": entry["ms_per_step"], "e2e": entry["e2e"], "gpu_launches": entry["gpu_launches"],
            "roofline": entry["roofline"], "clocks": entry["clocks"], "collective": entry["collective"],
            "peaks": ctx.peaks(),
        }
        if mgc is not None:
            line["multi_gpu_check"] = mgc
        if secondary:
            line["secondary"] = secondary
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_sample(main_spec, a.cpu_seconds)
            if not (a.tess or a.warp):
                try:
                    par = parity_sample((96, 96, 4, main_spec.bounces, main_spec.seed), r)
                    line["cpu_baseline"]["parity"] = par
                    line["divergent_ppm"] = par["divergent_ppm"]
                except Exception as ex:      # noqa: BLE001
                    line["cpu_baseline"]["parity"] = {"unavailable": f"{type(ex).__name__}: {ex}"[:200]}
                line["cpu_baseline"]["reference_python"] = python_reference_sample(4, 2)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_config1(a):
    """BASELINE.json configs[0]: objs/cornellroom.sdl as shipped (40x40, npaths 10 -> -r 10, -b 2): the unmodified Python
    reference on this host's cores next to the product on cuda:0."""
    import torch
    import pathtracerpython_b200 as ptb
    ref = python_reference_sample(10, 2)
    r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
    for _ in range(3):
        r.trace_paths(10, 2, seed=a.seed)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        acc, _ = r.trace_paths(10, 2, seed=a.seed)
        r.image(acc)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    # the same frame as a CUDA graph (ProgressiveFrame: ptb_trace_paths_at + tone map captured once, the sample offset advanced
    # on the device): a 40x40 frame is ~15 launches of microsecond kernels, i.e. launch-bound
    pf = ptb.ProgressiveFrame(r, 10, 2, seed=a.seed)
    for _ in range(5):
        pf.step()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(200):
        pf.step()
    e1.record()
    torch.cuda.synchronize()
    ms_graph = e0.elapsed_time(e1) / 200
    print(json.dumps({"metric": "Msamples/s", "value": 16000 / ms / 1e3, "unit": "Msamples/s", "n_gpus": 1, "ms_per_step": ms,
                      "config": {"workload": "cornell_box_40x40_10spp_b2 (objs/cornellroom.sdl as shipped)"},
                      "cuda_graph": {"value": 16000 / ms_graph / 1e3, "unit": "Msamples/s", "ms_per_step": ms_graph,
                                     "note": "the frame's launch sequence replayed as one CUDA graph, fresh samples every replay"},
                      "cpu_baseline": ref, "dtype": "f32"}))


if __name__ == "__main__":
    args = parse()
    if args.config1:
        run_config1(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
