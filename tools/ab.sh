timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['roofline']['stage_ms_per_step']; print(round(d['ms_per_step'],1), round(d['e2e']['seconds_per_step']*1e3,1), {k: round(v,1) for k,v in s.items() if v>1})"
