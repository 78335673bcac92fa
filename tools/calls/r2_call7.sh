timeout 600 python -m pytest tests/test_gpu_mcm.py -m gpu -q 2>&1 | grep -v Warning | tail -60
python bench.py --configs mcm --mcm-streams 2048 --no-ask --e2e-steps 0 --steps 3 --warmup 3 > gpurun_out/mcm_small.json 2> gpurun_out/mcm_small.err; echo rc=$?; tail -3 gpurun_out/mcm_small.err
python -c "
import json; d=json.load(open('gpurun_out/mcm_small.json')); print(json.dumps(d['configs'], indent=1))"
