timeout 600 python -m pytest tests/test_gpu_small_path.py -m gpu -x -q 2>&1 | grep -v Warning | tail -4
for n in 1 256 4096; do TMK_NO_CLUSTER=1 python tools/bench_links.py $n 100 2>&1 | tail -1; done
