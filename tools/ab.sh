timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for far in 0 1 2; do echo "BF_FAR=$far"; BF_FAR=$far BF_TIMING=1 timeout 300 python bench.py --steps 1 --warmup 0 --no-cpu-baseline --stars 33792 2>&1 >/dev/null | grep "bf timing" | head -3; done
