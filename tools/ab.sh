for e in 0 1; do if [ $e = 1 ]; then export BF_EXPERIMENT_NO_PDF_COPY=1; fi; BF_TAIL_BATCH=0 timeout 300 python bench.py --steps 3 --warmup 1 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('nocopy $e', round(d['ms_per_step'],1), round(d['e2e']['seconds_per_step']*1e3,1))"; done
