for v in _w1 _w2 ""; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  python bench.py --no-configs --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 6 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('VAR', '$v', round(d['value']), round(d['ms_per_step'],3), {k: round(x,3) for k,x in d['stage_ms_per_step'].items()})"
done
export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy_w2.so
python bench.py --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 3 --warmup 2 --configs man,chunk,c3,c4 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
for k,v in d['configs'].items(): print(k, round(v['msamples_per_s']), round(v['ms_per_step'],2), {a: round(b,2) for a,b in v.get('stage_ms',{}).items()}, v.get('parity'))"
