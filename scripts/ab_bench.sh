for a in 10 12 16; do
echo "AGG=$a"; if [ -z "$a" ]; then unset RNB_AGG_LEVELS; else export RNB_AGG_LEVELS=$a; fi; python bench.py --steps 32 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['stage_ms_per_launch'])"
done
