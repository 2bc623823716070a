timeout 300 python -m pytest tests/test_gpu_api.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
python - <<'PY'
import torch, time
x = torch.empty(1_500_000_000, dtype=torch.uint8, device="cuda")
h = torch.empty(1_500_000_000, dtype=torch.uint8, pin_memory=True)
for _ in range(2):
    torch.cuda.synchronize(); t = time.perf_counter(); h.copy_(x, non_blocking=True); torch.cuda.synchronize()
    print("D2H GB/s", 1.5 / (time.perf_counter() - t))
PY
timeout 500 python bench.py --steps 3 --warmup 1 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['seconds_per_step'], d['e2e']['value'])"
