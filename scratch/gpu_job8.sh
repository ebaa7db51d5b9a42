#!/bin/bash
# scratch/gpu_job8.sh - the 8-GPU batch (gpurun --gpus 8): multi-GPU correctness, the driver's scaling command at N = 8,
# and BASELINE.json configs[4] as stated (3840 x 2160, 4096 spp, 9 999 392 triangles, sample ranges over 8 GPUs)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
echo "== tests that need more than one GPU"
python -m pytest tests/test_gpu_multi.py tests/test_gpu_threads.py -m gpu -q 2>&1 | tail -4
$TR tests/multi_gpu_check.py 2>&1 | grep -v "^W\|OMP_NUM" | tail -4
echo "== bench.py --gpus 8 (headline + C5 strong / weak, multi_gpu_check in the line)"
$TR bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02ac_bench_8gpu.json 2> gpurun_out/r02ac_bench_8gpu.err; tail -c 400 gpurun_out/r02ac_bench_8gpu.err
echo "== configs[4] as stated: 4096 spp"
$TR bench.py --gpus 8 --tess 559 --width 3840 --height 2160 --spp 4096 --shard sample --steps 1 --warmup 1 --no-secondary --no-e2e-variants > gpurun_out/r02ac_bench_c5_8gpu_4096spp.json 2> gpurun_out/r02ac_c5.err; tail -c 300 gpurun_out/r02ac_c5.err
python - <<'PY'
import json
for f in ("r02ac_bench_8gpu", "r02ac_bench_c5_8gpu_4096spp"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, "value", round(d["value"], 2), "ms", round(d["ms_per_step"], 1), "e2e", d["e2e"].get("value"), d["e2e"].get("variant"), "check", d.get("multi_gpu_check"), "collective", d.get("collective"), d.get("cpu_baseline", {}).get("cores"))
        for s in d.get("secondary", []):
            print("  ", s["workload"], s["scaling"], round(s["value"], 1), "e2e", round(s["e2e"]["value"], 1), s["e2e"].get("variant"), {k: round(v["value"], 1) for k, v in s["e2e"].get("variants", {}).items()})
    except Exception as e:
        print(f, "FAILED", e)
PY
