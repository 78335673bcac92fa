timeout 900 python -m pytest tests/test_gpu_segwalk.py tests/test_gpu_configs.py -m gpu -x -q 2>&1 | grep -v Warning | tail -25
python bench.py --configs c4 --no-ask --e2e-steps 0 --steps 5 --warmup 3 > gpurun_out/c4_seg.json 2> gpurun_out/c4_seg.err; echo rc=$?; tail -3 gpurun_out/c4_seg.err
python -c "
import json; d=json.load(open('gpurun_out/c4_seg.json')); print(round(d['value']), d['stage_ms_per_step']); print(json.dumps(d['configs'], indent=1))"
