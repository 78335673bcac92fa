for v in "" _w2 _w4 _c4 _c6; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  python bench.py --no-configs --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 6 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('VAR', '$v', round(d['value']), round(d['ms_per_step'],3), {k: round(x,3) for k,x in d['stage_ms_per_step'].items()})"
done
