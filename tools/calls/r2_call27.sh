CMD="python bench.py --snrs=3 --streams 512 --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 1 --warmup 1"
$CMD 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print({k: round(v,3) for k,v in d['stage_ms_per_step'].items()}, d['decode']['walk_last_step'])"
ncu --set full --clock-control none --import-source on -k regex:"walk_kernel" -s 2 -c 1 -o gpurun_out/r02_walk_3db $CMD > gpurun_out/ncu_w3.log 2>&1; tail -1 gpurun_out/ncu_w3.log
