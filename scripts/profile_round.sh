#!/bin/bash
# Round profile pass (B200_PROFILING.md recipe): plain bench first, then the launch list, then one full capture
# of the kernels named on the command line (default: the fused field kernels and the marcher).
# Outputs under gpurun_out/ (scratch); summaries are copied into profiles/ by hand.
set -u
TAG=${1:-r01}
KERNELS=${2:-"regex:field_bwd_tc_kernel|field_fwd_tc_kernel|eng_march_kernel|eng_adam_kernel"}
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || { echo "plain bench failed"; tail -5 gpurun_out/${TAG}_bench.err; exit 1; }
cat gpurun_out/${TAG}_bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "$KERNELS" --launch-skip 8 --launch-count 8 \
  -o gpurun_out/${TAG}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/
