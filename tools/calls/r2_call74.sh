for v in "" _aw3 _aw4 _aw6 _aw8; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  python bench.py --no-configs --no-links --e2e-steps 0 --no-cpu --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('VAR $v', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'])"
done
