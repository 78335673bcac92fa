# per-SNR-group walk time and debug counters (512 streams each, as in C2), then C2 with the groups reversed
for s in -3 0 3 6 10 15 20 30; do
  python bench.py --snrs=$s --streams 512 --no-cpu --e2e-steps 0 --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('SNR', '$s', {k: round(v,3) for k,v in d['stage_ms_per_step'].items()}, d['decode']['walk_last_step'])"
done
python bench.py --snrs=30,20,15,10,6,3,0,-3 --no-cpu --e2e-steps 0 --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('REV', round(d['value']), d['stage_ms_per_step'])"
