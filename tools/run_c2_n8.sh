timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c2_n8b.json 2> gpurun_out/bench_c2_n8b.err; echo "c2 rc=$?"; tail -c 300 gpurun_out/bench_c2_n8b.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c2_n8b.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['seconds_per_step'], d['e2e']['host_cores_rank0'])"
nvidia-smi topo -m | head -14; lscpu | grep -i "numa\|socket\|^CPU(s)"
