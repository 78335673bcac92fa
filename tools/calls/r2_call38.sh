# order independence: the GPU suite with its files in reverse order, then the default order, in fresh processes
timeout 900 python -m pytest $(ls tests/test_gpu_*.py tests/test_zz_no_alloc.py | sort -r) -m gpu -x -q -p no:cacheprovider 2>&1 | grep -v Warning | tail -4
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -3
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
