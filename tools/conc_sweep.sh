for v in "TMK_SERIAL_CLUSTER=1" "TMK_CLUSTER_CTAS=1" "TMK_CLUSTER_CTAS=2" "TMK_CLUSTER_CTAS=3"; do
  env $v timeout 300 python bench.py --no-cpu --e2e-steps 0 --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('CONC', '$v', round(d['value']), round(d['ms_per_step'],2), d['stage_ms_per_step'], d['decode']['walk_last_step'])"
done
