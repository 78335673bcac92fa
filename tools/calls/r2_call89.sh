# lane-parallel exact replay: steps per lane and round (TMK_PAR_SEG) and the minimum worth a round (TMK_PAR_MIN) swept
for v in "" _ps256 _ps384 _ps768 _ps1024 _pm64 _pm256 ""; do
  export TMKPHY_LIB=$PWD/trackmaker-rs_b200/libtmkphy$v.so
  python bench.py --no-configs --no-links --no-ask --no-pcm16 --e2e-steps 0 --no-cpu --steps 6 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); s=d['stage_ms_per_step']; w=d['decode']['walk_last_step']; print('VAR $v', round(d['ms_per_step'],3), {k: round(x,3) for k,x in s.items()}, d['decode']['records'], w['replay_steps'], w['parallel_replay_steps'], w['parallel_replay_doubts'])"
done
