#!/bin/bash
# Parity + A/B of the marcher's empty-space test (coarse.cuh).  Usage: scripts/cull_check.sh
scripts/quick_check.sh
echo "--- RNB_NO_CULL=1"
RNB_NO_CULL=1 timeout 200 python bench.py --steps 32 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('rays/s', round(d['value']), 'ms/step', round(d['ms_per_step'],4)); print(d['roofline']['stage_ms_per_launch'])"
for cull in 0 1; do
  echo "--- render RNB_NO_CULL=$cull"
  RNB_NO_CULL=$cull RNB_RENDER_PROF=1 RNB_RENDER_SPR=32 timeout 300 python scripts/bench_render.py 2>&1 | grep -E "rnb_render|ms_per_frame|\{" | tail -12
done
