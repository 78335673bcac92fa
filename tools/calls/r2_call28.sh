python tools/bench_links.py 256 60 2>&1 | tail -1
python tools/bench_links.py 4096 60 2>&1 | tail -1
timeout 900 python -m pytest tests/test_gpu_links.py tests/test_gpu_parity.py tests/test_gpu_configs.py tests/test_gpu_segwalk.py -m gpu -x -q 2>&1 | grep -v Warning | tail -4
python bench.py --no-configs --no-ask --e2e-steps 2 --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(round(d['value']), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['stage_ms_per_step'].items()}, round(d['e2e']['value']), round(d['e2e_pcm16']['value']))"
