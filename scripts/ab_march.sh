# march stage time with / without chunk skipping on a workload: scripts/ab_march.sh <workload> "ENV=VAL" ...
W=$1; shift
for spec in "$@"; do
env $spec python bench.py --steps 20 --warmup 5 --no-cpu-baseline --workload $W 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['roofline']['stage_ms_per_launch']; print('$W [$spec]', 'step', round(d['ms_per_step'],4), 'value', round(d['value']), 'march', s['march'], 'bwd', s['field_bwd'], 'fwd', s['field_fwd'], 'spr', round(d['run']['samples_per_ray'],3), 'loss', round(d['final_loss'],4))"
done
