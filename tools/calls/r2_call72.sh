timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -2
python bench.py --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 6 --warmup 3 --configs c3,c5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(round(d['value']), round(d['ms_per_step'],3), {k: round(x,3) for k,x in d['stage_ms_per_step'].items()})
for k,v in d['configs'].items(): print(k, round(v['msamples_per_s']), round(v['ms_per_step'],2), {a: round(b,2) for a,b in v.get('stage_ms',{}).items()}, v.get('stage_ms_rank0'), v.get('parity'))"
