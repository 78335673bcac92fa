#!/usr/bin/env python
"""Per-source-line summary of an ncu report: share of stall samples and of executed instructions.
usage: ncu_lines.py report.ncu-rep kernel_regex [top_n]   (runs `ncu -i ... --page source --print-source cuda,sass`)"""
import csv
import io
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
    rows, fname, hdr = [], "", None
    for x in csv.reader(io.StringIO(out)):
        if len(x) == 2 and x[0] == "File Path":
            fname = x[1].split("/")[-1]
        elif len(x) > 8 and x[0] == "Line No":
            hdr = x
        elif hdr and len(x) == len(hdr) and x[0].isdigit():
            d = dict(zip(hdr, x))
            stalls = {k[6:]: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v)}
            rows.append((int(d["# Samples"]), int(d["Instructions Executed"]), fname, int(x[0]), x[1].strip(), stalls))
    ts, ti = sum(r[0] for r in rows) or 1, sum(r[1] for r in rows) or 1
    print(f"kernel {kern}: {ts} samples, {ti} warp instructions")
    for s, i, f, ln, src, st in sorted(rows, key=lambda r: (r[0], r[1]), reverse=True)[:top]:
        tops = ",".join(f"{k}:{v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{s / ts * 100:5.1f}% smp {i / ti * 100:5.1f}% ins {f}:{ln:<4} {src[:90]}  [{tops}]")


if __name__ == "__main__":
    main()
