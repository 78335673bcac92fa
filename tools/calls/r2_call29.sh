timeout 600 python -m pytest tests/test_gpu_ingest.py -m gpu -x -q 2>&1 | grep -v Warning | tail -6
python bench.py --impl reference --steps 2 --warmup 1 | cut -c1-400
