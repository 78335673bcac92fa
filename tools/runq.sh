# quick A/B: scan/filter tests, then the bench without the cpu / e2e legs
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('RES', round(d['value']), {k: round(v,3) for k,v in d['stage_ms_per_step'].items()}, round(d['roofline']['frac'],4))"; done
