timeout 600 python -m pytest tests/test_gpu_links.py tests/test_gpu_small_path.py tests/test_gpu_reasm.py -m gpu -x -q 2>&1 | grep -v Warning | tail -2
for n in 1 256; do python tools/bench_links.py $n 100 2>&1 | tail -1; done
