for v in "" _k64 _k512; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  timeout 300 python bench.py --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('LIB', '$v', round(d['value']), round(d['ms_per_step'],2), {k: round(x,2) for k,x in d['stage_ms_per_step'].items()}, d['decode']['walk_last_step'])"
  for s in 10 30; do timeout 300 python bench.py --snrs=$s --streams 512 --no-cpu --e2e-steps 0 --no-ask --no-configs --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('  snr $s', '$v', {k: round(x,2) for k,x in d['stage_ms_per_step'].items()}, d['decode']['walk_last_step']['blocks_examined'])"; done
done
