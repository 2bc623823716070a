timeout 1500 python bench.py --config C4 --pdf2d --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "c4 rc=$?"; tail -c 300 gpurun_out/bench_c4.err
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --stars 33792 > gpurun_out/pre_list.json 2>/dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --stars 33792 > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
