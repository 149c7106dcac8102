#!/bin/bash
for p in 1 0; do
  RNB_BENCH_PRIO=$p timeout 200 python bench.py --steps 48 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('prio=$p rays/s', round(d['value']), 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']))"
done
