#!/bin/bash
# One full ncu capture (with source correlation) of the backward field kernel from a short bench run -> gpurun_out/$1.ncu-rep
OUT=${1:-bwd}
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:field_bwd_tc_kernel" --launch-skip 4 --launch-count 1 \
  -o gpurun_out/${OUT} -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${OUT}_ncu.log 2>&1
echo "ncu rc=$?"
