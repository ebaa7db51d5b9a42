#!/bin/bash
# scratch/gpu_job.sh - the batch a gpurun slot executes (results under gpurun_out/)
mkdir -p gpurun_out
echo "== full GPU test suite (final build)"
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/pytest_r02d.log; tail -3 gpurun_out/pytest_r02d.log
echo "== default bench line (headline + secondary), reference arm, config 1"
python bench.py --steps 5 --warmup 3 > gpurun_out/r02d_bench_1gpu.json 2> gpurun_out/r02d_bench_1gpu.err; tail -c 300 gpurun_out/r02d_bench_1gpu.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02d_bench_reference_arm.json 2> gpurun_out/r02d_ref.err
python bench.py --config1 > gpurun_out/r02d_bench_config1.json 2> gpurun_out/r02d_c1.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02d_bench_1gpu.json"))
print("headline", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 3), "ppm", d.get("divergent_ppm"))
for s in d.get("secondary", []):
    print(s["workload"], round(s["value"], 1), "e2e", round(s["e2e"]["value"], 1), "frac", round(s["roofline"]["frac"], 3), s["roofline"]["kernel_ms"])
print(open("gpurun_out/r02d_bench_config1.json").read()[:600])
print(open("gpurun_out/r02d_bench_reference_arm.json").read()[:300])
PY
echo "== ncu: LBVH launch list (1 M triangles, 4 spp) + full captures of both traversal kernels"
CMD="python bench.py --tess 177 --spp 4 --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
$CMD > gpurun_out/r02d_plain_tess177_spp4.json 2> gpurun_out/r02d_plain_tess177_spp4.err &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r02d_launches_tess177_spp4.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_closest_bvh32_dyn -s 4 -c 2 -f -o gpurun_out/r02d_closest $CMD > gpurun_out/ncu2.log 2>&1
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_shadow_bvh32_dyn -s 3 -c 2 -f -o gpurun_out/r02d_shadow $CMD > gpurun_out/ncu3.log 2>&1
for k in closest shadow; do
  ncu -i gpurun_out/r02d_$k.ncu-rep --page raw --csv > gpurun_out/r02d_ncu_full_raw_$k.csv 2>/dev/null
  rm -f gpurun_out/r02d_$k.ncu-rep
done
ls -la gpurun_out | grep r02d
