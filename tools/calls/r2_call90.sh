# lane-parallel exact replay: sample loads two iterations ahead of the chain (TMK_PAR_DEPTH=2) against one; C2 (2-CTA walk) and
# the same samples as 16 384 shorter captures (3-CTA walk)
for shape in "4096 480000" "16384 120000"; do
 set -- $shape
 for v in "" _pd2 ""; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  python bench.py --streams $1 --samples $2 --no-configs --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 6 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); s=d['stage_ms_per_step']; w=d['decode']['walk_last_step']; print('VAR $v $1x$2', round(d['ms_per_step'],3), {k: round(x,3) for k,x in s.items()}, d['decode']['records'], w['replay_steps'], w['parallel_replay_steps'])"
 done
done
