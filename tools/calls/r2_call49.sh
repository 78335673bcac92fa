python bench.py --no-configs --no-ask --no-pcm16 --e2e-steps 1 --steps 3 --cpu-max-captures 64 2>gpurun_out/b49.err | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(json.dumps(d['links'], indent=1)); print(d['value'], d['ms_per_step'])"
tail -3 gpurun_out/b49.err
