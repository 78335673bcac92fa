for n in 1 256 4096; do TMK_NO_CLUSTER=1 python tools/bench_links.py $n 100; done
