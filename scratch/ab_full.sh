#!/bin/bash
run() { python bench.py --tess ${TESS:-177} --spp ${SPP:-64} ${EXTRA} --steps 3 --warmup 3 --no-secondary --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().strip().splitlines() if l.startswith('{')][-1]); k=d['roofline']['kernel_ms']; print('$1', round(d['value'],1), 'ms/step', round(d['ms_per_step'],2), 'sum', round(sum(k.values())/3,2), k)"; }
for v in "$@"; do if [ "$v" = cur ]; then run cur; else PTB_LIB_PATH=scratch/libptb_$v.so run $v; fi; done
