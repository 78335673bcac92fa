for m in 100000 512 256 128 64 32; do
  TMK_RESIDENT_CHUNK_MSAMPLES=$m python bench.py --no-cpu --e2e-steps 0 --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('CHUNK', '$m', round(d['value']), d['ms_per_step'], d['stage_ms_per_step'])"
done
