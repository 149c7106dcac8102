#!/bin/bash
# Parity tests that touch the hot kernels + a short bench; prints the stage table.  Usage: scripts/quick_check.sh [pytest targets]
T=${@:-tests/test_golden.py tests/test_vren_gpu.py tests/test_field_gpu.py tests/test_engine_gpu.py tests/test_render_gpu.py tests/test_grid_gpu.py}
timeout 900 python -m pytest $T -x -q 2>&1 | tail -8
timeout 200 python bench.py --steps 32 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('rays/s', round(d['value']), 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']), 'spr', d['config']['samples_per_ray'])
print(d['roofline']['stage_ms_per_launch'])"
