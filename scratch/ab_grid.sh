#!/bin/bash
for g in "$@"; do PTB_FUSED_GRID=$g python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().strip().splitlines() if l.startswith('{')][-1]); k=d['roofline']['kernel_ms']; print('grid x$g', round(d['value'],1), 'ms/frame', round(d['ms_per_step'],2), 'fused ms', k.get('bounce_fused'))"; done
