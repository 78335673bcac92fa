timeout 900 python -m pytest tests/test_gpu_ask.py tests/test_gpu_mcm.py tests/test_gpu_configs.py tests/test_gpu_ingest.py -m gpu -x -q 2>&1 | grep -v Warning | tail -12
python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(json.dumps(d['chirp_bpsk']))"
