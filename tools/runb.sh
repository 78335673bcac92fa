timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 5 --warmup 3 > gpurun_out/b3.json 2> gpurun_out/b3.err; echo rc=$?; tail -3 gpurun_out/b3.err
python -c "
import json
d=json.load(open('gpurun_out/b3.json'))
print('RES', round(d['value']), round(d['e2e']['value']) if d.get('e2e') else None, {k: round(v,3) for k,v in d['stage_ms_per_step'].items()}, d['cpu_baseline']['parity_vs_gpu'], d['decode']['walk_last_step'])"
