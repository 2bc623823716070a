for c in 96 192; do BF_DENSE_WAVES=$c timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['roofline']['stage_ms_per_step']; print('waves $c', round(d['ms_per_step'],1), {k: round(v,1) for k,v in s.items() if k in ('presort','dense','select')})"; done
