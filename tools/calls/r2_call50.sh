timeout 600 python -m pytest tests/test_gpu_ask.py tests/test_gpu_mcm.py -m gpu -x -q 2>&1 | grep -v Warning | tail -2
python bench.py --no-configs --no-links --e2e-steps 0 --no-cpu --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('chirp', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'])"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ask_ -c 30 --csv --log-file gpurun_out/ask50.csv python bench.py --no-configs --no-links --e2e-steps 0 --no-cpu --steps 1 --warmup 1 > /dev/null 2>&1
grep -v "^==" gpurun_out/ask50.csv | awk -F'","' 'NR>1{print substr($5,1,40), $NF}' | sort | uniq -c | sort -k4 -n | tail -8
