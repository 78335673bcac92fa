timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -2
for n in 1 256 4096; do python tools/bench_links.py $n 100; done
TMK_SYNC_SPIN=1 python tools/bench_links.py 256 100
TMK_SYNC_SPIN=1 python tools/bench_links.py 1 100
