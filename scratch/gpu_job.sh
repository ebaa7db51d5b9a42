#!/bin/bash
mkdir -p gpurun_out
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    r = d["roofline"]
    print(sys.argv[1], "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1), "e2e ms", round(d["e2e"]["ms_per_step"], 1),
          "kernel_ms", r["kernel_ms"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
echo "== C5 per GPU: end-to-end step after raising the pool's keep threshold"
PTB_TIMING=1 python bench.py --tess 559 --width 3840 --height 2160 --spp 8 --shard sample --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/r02f_c5.json 2> gpurun_out/r02f_c5.err
summ gpurun_out/r02f_c5.json; grep "bench e2e" gpurun_out/r02f_c5.err | tail -3
echo "== C4 64 spp: L2 persistence window over the node arrays (MB set aside)"
for mb in 0 48 90; do
  PTB_L2_PERSIST_MB=$mb python bench.py --tess 177 --spp 64 --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/r02f_c4_l2_$mb.json 2> gpurun_out/r02f_c4_l2_$mb.err
  summ gpurun_out/r02f_c4_l2_$mb.json
done
echo "== C5 8 spp with the window"
PTB_L2_PERSIST_MB=90 python bench.py --tess 559 --width 3840 --height 2160 --spp 8 --shard sample --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/r02f_c5_l2_90.json 2> gpurun_out/r02f_c5_l2_90.err
summ gpurun_out/r02f_c5_l2_90.json
