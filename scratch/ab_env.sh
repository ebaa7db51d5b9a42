#!/bin/bash
# A/B of an environment switch on the 1M-triangle scene: Msamples/s, per-kernel ms, nodes per ray
run() { python bench.py --tess ${TESS:-177} --spp ${SPP:-32} --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms']; print('$1', round(d['value'],1), {x:k.get(x) for x in ('closest_hit','shadow','shade','ray_sort')}, {a:round(b,2) for a,b in d['roofline']['per_ray'].items()})"; }
run default
env $1=1 bash -c "$(declare -f run); run $1"
