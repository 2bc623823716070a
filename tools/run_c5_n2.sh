timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --config C5 --stars 65536 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c5_n2.json 2> gpurun_out/bench_c5_n2.err; echo "c5 rc=$?"; tail -c 300 gpurun_out/bench_c5_n2.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c5_n2.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d.get('collective'), d['roofline'].get('stage_ms_per_step'))"
timeout 600 python -m pytest tests/test_subgrid_gpu.py -m gpu -x -q 2>&1 | tail -2
