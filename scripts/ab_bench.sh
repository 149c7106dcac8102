timeout 240 python -m pytest tests/test_field_gpu.py tests/test_engine_gpu.py -x -q 2>&1 | tail -2
for a in "" 0 3 8; do
echo "AGG=$a"; if [ -z "$a" ]; then unset RNB_AGG_LEVELS; else export RNB_AGG_LEVELS=$a; fi; python bench.py --steps 32 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['stage_ms_per_launch'])"
done
