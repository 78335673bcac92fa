for v in "" _c5 _c6; do
  TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so timeout 300 python bench.py --no-cpu --e2e-steps 0 --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('LIB', '$v', round(d['value']), round(d['ms_per_step'],2), d['stage_ms_per_step'], d['decode']['walk_last_step']['cluster_us'])"
done
