#!/bin/bash
# Parity + A/B of the marcher's empty-space test (coarse.cuh).  Usage: scripts/cull_check.sh [workload]
W=${1:-radnerf_360_k4}
timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_render_gpu.py tests/test_golden.py -x -q 2>&1 | tail -4
for cull in 0 1; do
  echo "--- $W RNB_NO_CULL=$cull"
  RNB_NO_CULL=$cull timeout 300 python bench.py --workload $W --steps 32 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('rays/s', round(d['value']), 'ms/step', round(d['ms_per_step'],4), 'spr', d['config'].get('samples_per_ray')); print(d['roofline']['stage_ms_per_launch'])"
done
