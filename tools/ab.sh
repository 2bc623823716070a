show() { python -c "import json,sys; d=json.loads(sys.stdin.read()); s=d['roofline']['stage_ms_per_step']; print('$1', round(d['ms_per_step'],1), {k: round(v,1) for k,v in s.items() if v>1})"; }
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline | show default
for b in -1 4 5; do BF_SORT_BAND=$b timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline | show band$b; done
