timeout 900 python -m pytest tests/test_gpu_ask.py tests/test_gpu_mcm.py tests/test_gpu_configs.py tests/test_gpu_ingest.py -m gpu -x -q 2>&1 | grep -v Warning | tail -6
for c in 1 2 4; do
  export TMK_ASK_CHUNKS=$c
  python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('CHUNKS $c', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'])"
done
unset TMK_ASK_CHUNKS
python bench.py --configs c5,mcm --no-ask --e2e-steps 0 --no-cpu --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
for k,v in d['configs'].items(): print(k, round(v['msamples_per_s']), round(v['ms_per_step'],2), v.get('stage_ms_rank0'), {m:(round(x['msamples_per_s']), x['stage_ms']) for m,x in v.get('modes',{}).items()})"
