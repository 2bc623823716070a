set -e
timeout 300 python bench.py --steps 1 --warmup 0 --no-cpu-baseline --stars 33792 > gpurun_out/pre_ncu.json 2> gpurun_out/pre_ncu.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dense_chi2_kernel -s 1 -c 1 -o gpurun_out/r02_dense_pruned -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --stars 33792 > gpurun_out/ncu_dense.log 2>&1
