timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 3 --warmup 2 --no-cpu --e2e-steps 0 --no-ask --configs c4 > gpurun_out/r2_b3.json 2> gpurun_out/r2_b3.err; echo rc=$?; tail -3 gpurun_out/r2_b3.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_b3.json'))
for k,v in d.get('configs',{}).items(): print(k, json.dumps(v))
PY
CMD="python bench.py --snrs=3 --streams 512 --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 1 --warmup 1"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:walk_kernel -s 1 -c 1 -o gpurun_out/r2_walk3db $CMD > gpurun_out/ncu.log 2>&1; tail -2 gpurun_out/ncu.log
