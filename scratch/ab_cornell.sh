#!/bin/bash
# A/B of library variants on the headline workload: name, Msamples/s, fused-kernel ms.  usage: ab_cornell.sh name ...   (scratch/libptb_<name>.so; "cur" = the in-tree library)
run() { python bench.py --steps ${STEPS:-5} --warmup 3 --no-secondary --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().strip().splitlines() if l.startswith('{')][-1]); k=d['roofline']['kernel_ms']; print('$1', round(d['value'],1), 'ms/frame', round(d['ms_per_step'],2), 'fused ms', k.get('bounce_fused'), 'clk', d['clocks']['sm_mhz'])"; }
for v in "$@"; do if [ "$v" = cur ]; then run cur; else PTB_LIB_PATH=scratch/libptb_$v.so run $v; fi; done
