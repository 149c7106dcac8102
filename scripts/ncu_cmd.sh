#!/bin/bash
# scripts/ncu_cmd.sh <kernel-regex> <out-tag> <launch-skip> <command...>: plain run first, then one --set full capture
K=$1; OUT=$2; SKIP=$3; shift 3
timeout 300 "$@" > gpurun_out/${OUT}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${OUT}_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$K" --launch-skip $SKIP --launch-count 2 \
  -o gpurun_out/${OUT} -f "$@" > gpurun_out/${OUT}_ncu.log 2>&1
echo "ncu rc=$?"
