#!/bin/bash
# A/B of the table-gradient aggregation knobs.  Usage: scripts/ab_agg360.sh workload "L R" "L R" ...
W=$1; shift
for cfg in "$@"; do
  set -- $cfg
  [ "$1" = "-" ] && unset RNB_AGG_LEVELS || export RNB_AGG_LEVELS=$1
  [ "$2" = "-" ] && unset RNB_AGG_LONG_RES || export RNB_AGG_LONG_RES=$2
  echo -n "$W AGG_LEVELS=$1 LONG_RES=$2: "
  timeout 400 python bench.py --workload $W --steps 16 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), d['roofline']['stage_ms_per_launch']['field_bwd'])"
done
