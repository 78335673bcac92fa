timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_b2.json 2> gpurun_out/r2_b2.err; echo rc=$?; tail -5 gpurun_out/r2_b2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_b2.json'))
print('RES', round(d['value']), round(d['e2e']['value']), d['stage_ms_per_step'], d['roofline']['frac'], d['cpu_baseline'])
for k,v in d.get('configs',{}).items(): print(k, json.dumps(v))
print('ASK', json.dumps(d.get('chirp_bpsk')))
PY
