#!/bin/bash
# A/B of library variants scratch/libptb_<name>.so on the 1M-triangle scene: name, Msamples/s, per-kernel ms
run() { python bench.py --tess ${TESS:-177} --spp ${SPP:-32} --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms']; print('$1', round(d['value'],1), {x:k.get(x) for x in ('closest_hit','shadow','shade','ray_sort')})"; }
run base
for v in "$@"; do PTB_LIB_PATH=scratch/libptb_$v.so run $v; done
