run() { env "$@" python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['roofline']['stage_ms_per_launch']; print('$*', 'step', round(d['ms_per_step'],4), 'noflush', round(d['ms_per_step_noflush'],4), 'bwd', s['field_bwd'], 'fwd', s['field_fwd'], 'adam', s['adam'], 'loss', d['final_loss'])"; }
run RNB_AGG_HALF=0
run RNB_AGG_HALF=1
run RNB_AGG_HALF=1 RNB_AGG_LEVELS=10
run RNB_AGG_HALF=1 RNB_AGG_LEVELS=12
