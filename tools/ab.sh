timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -m gpu -x -q 2>&1 | tail -2
BF_TIMING=1 timeout 300 python bench.py --steps 1 --warmup 0 --no-cpu-baseline --stars 33792 2>&1 >/dev/null | grep "bf timing" | head -3
