#!/bin/bash
# Multi-GPU measurements of one box: training (weak scaling) and sharded frame rendering.  Output: gpurun_out/scale_*.json
# Usage: scripts/scale_run.sh "N workload" ...   (default: the round's standard set on an 8-GPU box)
mkdir -p gpurun_out
run() {  # N workload
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 295$1$1 \
    bench.py --gpus $1 --steps 48 --warmup 5 --no-cpu-baseline --workload $2 2> gpurun_out/scale_$2_n$1.err | tail -1 > gpurun_out/scale_$2_n$1.json
  python - <<PY
import json
d=json.load(open("gpurun_out/scale_$2_n$1.json"))
print("$2 N=$1 rays/s", round(d["value"]), "ms/step", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), d["clocks"]["reasons"])
PY
}
if [ $# -gt 0 ]; then
  for spec in "$@"; do run $spec; done
  exit 0
fi
run 8 radnerf_tat_k2
run 4 radnerf_tat_k2
run 2 radnerf_tat_k2
run 8 radnerf_360_k4
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29577 \
  scripts/bench_render_dist.py radnerf_tat_k8_infer 2> gpurun_out/scale_render_n8.err | tail -1 | tee gpurun_out/scale_render_n8.json
