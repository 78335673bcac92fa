TMK_TRACE_HOST=1 python tools/bench_links.py 256 40 2>&1 | tail -12
TMK_TRACE_HOST=1 python tools/bench_links.py 1 40 2>&1 | tail -5
