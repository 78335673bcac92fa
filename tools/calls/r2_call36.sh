timeout 600 python -m pytest tests/test_gpu_ask.py -m gpu -x -q 2>&1 | grep -v Warning | tail -3
python bench.py --configs c5 --no-ask --e2e-steps 0 --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
for k,v in d['configs'].items(): print(k, round(v['msamples_per_s']), round(v['ms_per_step'],2), v.get('stage_ms_rank0'), v.get('parity'))"
