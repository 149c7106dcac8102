#!/bin/bash
# One full ncu capture of the kernels matching $1 (regex) from a short bench run; report to gpurun_out/$2.ncu-rep
set -u
K=${1:-field_bwd_tc_kernel}
OUT=${2:-cap}
timeout 200 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/${OUT}_plain.json 2> gpurun_out/${OUT}_plain.err || { echo "plain run failed"; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$K" --launch-skip 4 --launch-count 2 \
  -o gpurun_out/${OUT} -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${OUT}_ncu.log 2>&1
echo "ncu rc=$?"
