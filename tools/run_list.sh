timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --stars 33792 > gpurun_out/pre_list.json 2>/dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --stars 33792 > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/pre_list.json').read().strip().splitlines()[-1]); print(d['ms_per_step'])"
