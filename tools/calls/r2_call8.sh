timeout 900 python -m pytest tests/test_gpu_ask.py tests/test_gpu_mcm.py tests/test_gpu_configs.py -m gpu -q 2>&1 | grep -v Warning | tail -8
for v in 1 0; do
  if [ $v = 1 ]; then export TMK_ASK_NO_CHK=1; else unset TMK_ASK_NO_CHK; fi
  python bench.py --no-configs --no-cpu --e2e-steps 0 --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('NO_CHK=$v', json.dumps(d['chirp_bpsk']))"
done
