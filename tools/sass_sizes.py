#!/usr/bin/env python
"""Instruction count per function (kernels and their non-inlined device functions) of an object file.
usage: sass_sizes.py file.o"""
import re, subprocess, sys, tempfile, os, glob
o = os.path.abspath(sys.argv[1])
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", o], cwd=d, stdout=subprocess.DEVNULL)
    for cubin in glob.glob(d + "/*.cubin"):
        out = subprocess.run(["nvdisasm", "-c", cubin], capture_output=True, text=True).stdout
        name, n, rows = None, 0, []
        for line in out.splitlines():
            m = re.match(r"^([$_.A-Za-z0-9]+):\s*$", line)
            if m and not m.group(1).startswith(".L_"):
                if name: rows.append((n, name))
                name, n = m.group(1), 0
            elif re.match(r"^\s+/\*[0-9a-f]+\*/\s+\S", line):
                n += 1
        if name: rows.append((n, name))
        for n, name in sorted(rows):
            if n: print(f"{n:7d}  {name.split('$')[-1][-70:]}")
