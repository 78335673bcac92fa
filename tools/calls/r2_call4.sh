timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 5 --warmup 3 --no-configs > gpurun_out/r2_b4.json 2> gpurun_out/r2_b4.err; echo rc=$?; tail -3 gpurun_out/r2_b4.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_b4.json'))
print('RES', round(d['value']), round(d['e2e']['value']), d['stage_ms_per_step'], d['roofline']['frac'], d['cpu_baseline']['parity_vs_gpu'], d['decode']['walk_last_step'])
PY
for s in 3 10; do python bench.py --snrs=$s --streams 512 --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('SNR', '$s', {k: round(v,3) for k,v in d['stage_ms_per_step'].items()}, d['decode']['walk_last_step'])"; done
