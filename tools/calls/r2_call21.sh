for t in 6216 9000 12500 17000 25000; do
  export TMK_ASK_SEGMENTS=$t
  python bench.py --no-configs --e2e-steps 0 --no-cpu --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['chirp_bpsk']; print('ASK_SEGMENTS $t', round(c['value']), round(c['filter_ms'],2), round(c['walk_ms'],2), c['frames'])"
done
unset TMK_ASK_SEGMENTS
python tools/bench_links.py 2>&1 | tail -4
