#!/bin/bash
# A/B of one environment switch on the default bench: scripts/ab_env.sh VAR "v1 v2 ..." [bench args]
VAR=$1; VALS=$2; shift 2
for v in $VALS; do
  for rep in 1 2; do
    env $VAR=$v timeout 200 python bench.py --no-cpu-baseline "$@" 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$VAR=$v', 'ms/step', round(d['ms_per_step'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4), 'rays/s', round(d['value']))"
  done
done
