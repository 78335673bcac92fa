timeout 600 python -m pytest tests/test_gpu_ask.py -m gpu -x -q 2>&1 | grep -v Warning | tail -3
python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('kSeg 8192', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'])"
