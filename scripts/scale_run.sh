#!/bin/bash
# Multi-GPU measurements of one box.  Output: gpurun_out/scale_<workload>_n<N>_<scaling>.json
# Every run is bounded by `timeout 150` (a hung multi-GPU run is charged N x its wall time) and gets its own rendezvous port.
# Usage: scripts/scale_run.sh "N workload [weak|strong] [ENV=VAL ...]" ...
mkdir -p gpurun_out
run() {  # N workload scaling env...
  local N=$1 W=$2 S=${3:-weak}; shift 3 2>/dev/null || shift $#
  local tag=scale_${W}_n${N}_${S}$(echo "$*" | tr -c 'A-Za-z0-9=\n' '_' | sed 's/^_*$//;s/^/_/;s/^_$//')
  if [ "$N" = 1 ]; then
    env "$@" timeout 150 python bench.py --gpus 1 --steps 48 --warmup 5 --no-cpu-baseline --workload $W --scaling $S 2> gpurun_out/$tag.err | tail -1 > gpurun_out/$tag.json
  else
    PORT=$((29500 + RANDOM % 400)); env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --steps 48 --warmup 5 --no-cpu-baseline --workload $W --scaling $S 2> gpurun_out/$tag.err | tail -1 > gpurun_out/$tag.json
  fi
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/$tag.json"))
    s=d["roofline"]["stage_ms_per_launch"]
    print("$tag rays/s", round(d["value"]), "ms/step", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), "replica", d.get("replica_check"), d["clocks"]["reasons"],
          {k: s[k] for k in ("march","field_fwd","field_bwd","adam","barrier") if k in s})
except Exception as e:
    print("$tag FAILED", e); print(open("gpurun_out/$tag.err").read()[-600:])
PY
}
for spec in "$@"; do run $spec; done
