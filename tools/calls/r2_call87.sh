# cluster pass: CTAs per SM re-swept on the round-2 kernel (the round-1 sweep predates its restructuring)
for v in "" _cc4 _cc6; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  python bench.py --no-configs --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 6 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); s=d['stage_ms_per_step']; print('VAR $v', round(d['ms_per_step'],3), {k: round(x,3) for k,x in s.items()}, d['decode']['records'])"
done
