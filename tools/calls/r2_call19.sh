python tools/ask_split_probe.py 2>&1 | tail -6
