timeout 300 python -m pytest tests/test_gpu_api.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
timeout 500 python bench.py --steps 3 --warmup 1 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['seconds_per_step'], d['e2e']['value'])"
