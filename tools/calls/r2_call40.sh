timeout 600 python -m pytest tests/test_gpu_ask.py tests/test_gpu_mcm.py -m gpu -x -q 2>&1 | grep -v Warning | tail -3
python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('power bulk', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'], c.get('parity'))"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ask_ -c 40 --csv --log-file gpurun_out/ask_launch40.csv python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 1 --warmup 1 > /dev/null 2>&1
python tools/ncu_sum.py gpurun_out/ask_launch40.csv 2>/dev/null | head -12
