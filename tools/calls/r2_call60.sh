timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | grep -v Warning | tail -3
python tools/bench_links.py 256 100
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
