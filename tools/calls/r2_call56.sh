timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -3
python bench.py --no-configs --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(round(d['value']), round(d['ms_per_step'],3), {k: round(x,3) for k,x in d['stage_ms_per_step'].items()}, d['roofline']['kernels'] if 'kernels' in d['roofline'] else '')"
