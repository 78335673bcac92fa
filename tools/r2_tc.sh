timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for i in 1 2; do python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 0 --no-ask --no-configs 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('RES', round(d['value']), {k: round(v,3) for k,v in d['stage_ms_per_step'].items()}, round(d['roofline']['frac'],4), d['decode']['walk_last_step'])"; done
for s in 0 3 6 10 30; do python bench.py --snrs=$s --streams 512 --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('SNR', '$s', {k: round(v,3) for k,v in d['stage_ms_per_step'].items()})"; done
