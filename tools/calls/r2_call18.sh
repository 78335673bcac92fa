for cfg in "0 131072" "8192 131072" "12288 131072" "16384 100000" "32768 65536" "32768 32768"; do
  set -- $cfg
  export TMK_WALK_SEGMENTS=$1 TMK_WALK_SEG_MIN=$2
  python bench.py --no-configs --no-ask --e2e-steps 0 --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('SEG $1 MIN $2', round(d['value']), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['stage_ms_per_step'].items()}, d['decode']['records'])"
done
