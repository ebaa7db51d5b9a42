#!/bin/bash
for g in "$@"; do PTB_SHADE_GRID_ENV=$g python bench.py --tess 177 --spp 64 --steps 3 --warmup 3 --no-secondary --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().strip().splitlines() if l.startswith('{')][-1]); k=d['roofline']['kernel_ms']; print('shade grid x$g', round(d['value'],1), 'sum', round(sum(k.values())/3,2), {x:k.get(x) for x in ('closest_hit','shadow','shade','ray_sort')})"; done
