timeout 600 python -m pytest tests/test_gpu_reasm.py tests/test_gpu_pcm_frag.py -m gpu -x -q 2>&1 | grep -v Warning | tail -15
