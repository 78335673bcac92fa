timeout 900 python -m pytest tests/test_gpu_small_path.py tests/test_gpu_links.py -m gpu -x -q 2>&1 | grep -v Warning | tail -8
