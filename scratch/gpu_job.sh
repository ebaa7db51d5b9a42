#!/bin/bash
# scratch/gpu_job.sh - the batch a gpurun slot executes (A/B builds live under scratch/ab/, results under gpurun_out/)
mkdir -p gpurun_out
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    r = d["roofline"]
    print(sys.argv[1], "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1),
          "kernel_ms", r["kernel_ms"], "per_ray", r.get("per_ray"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
echo "== Cornell headline: blocks per SM of the packed kernel (base = 4, 64 registers)"
for v in base mb3 mb5 lim; do
  PTB_LIB_PATH=$PWD/scratch/ab/libptb_$v.so python bench.py --steps 4 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_cornell_$v.json 2> gpurun_out/bench_cornell_$v.err
  summ gpurun_out/bench_cornell_$v.json
done
echo "== LBVH 1 M triangles, 64 spp: base vs pending-leaf slot in the closest-hit kernel"
for v in base defer top64 top256 top256defer; do
  PTB_LIB_PATH=$PWD/scratch/ab/libptb_$v.so python bench.py --tess 177 --spp 64 --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_c4_$v.json 2> gpurun_out/bench_c4_$v.err
  summ gpurun_out/bench_c4_$v.json
done
echo "== LBVH tests with the pending-leaf build, then with the staged top + pending leaf"
PTB_LIB_PATH=$PWD/scratch/ab/libptb_defer.so python -m pytest tests/test_gpu_bvh.py -m gpu -q -k "not million and not file_loader" 2>&1 | tail -5
PTB_LIB_PATH=$PWD/scratch/ab/libptb_top256defer.so python -m pytest tests/test_gpu_bvh.py tests/test_gpu_physical.py tests/test_gpu_soup.py -m gpu -q -k "not ten_million and not file_loader" 2>&1 | tail -8
echo "== ncu: Cornell 64 spp launch list + full capture of the packed fused kernel"
CMD="python bench.py --spp 64 --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
$CMD > gpurun_out/r02c_plain_cornell_spp64.json 2> gpurun_out/r02c_plain_cornell_spp64.err &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r02c_launches_cornell_spp64.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_bounce_planes32x2 -s 3 -c 2 -f -o gpurun_out/r02c_fused_x2 $CMD > gpurun_out/ncu2.log 2>&1
ncu -i gpurun_out/r02c_fused_x2.ncu-rep --page raw --csv > gpurun_out/r02c_ncu_full_raw_fused_x2.csv 2>/dev/null
ncu -i gpurun_out/r02c_fused_x2.ncu-rep --page source --csv > gpurun_out/r02c_ncu_source_fused_x2.csv 2>/dev/null
ls -la gpurun_out | tail -12
