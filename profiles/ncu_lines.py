#!/usr/bin/env python
"""Per-CUDA-source-line instruction counts from an .ncu-rep captured with --import-source on (needs -lineinfo)."""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = ""
lines = []
hdr = None
for row in csv.reader(io.StringIO(raw)):
    if not row:
        continue
    if row[0] == "File Path":
        cur_file = row[1].split("/")[-1]
    elif row[0] == "Line No":
        hdr = {k: i for i, k in enumerate(row)}
    elif hdr and len(row) > 8 and row[2] == "-":
        try:
            n = int(row[hdr["Instructions Executed"]] or 0)
            smp = int(row[4] or 0)
        except ValueError:
            continue
        lines.append((n, smp, cur_file, row[0], row[1].strip()[:90], row[hdr["Avg. Threads Executed"]]))
tot = sum(l[0] for l in lines) or 1
tots = sum(l[1] for l in lines) or 1
print("total warp instructions attributed to source lines: %d" % tot)
for n, smp, f, ln, src, thr in sorted(lines, reverse=True)[:top]:
    print("%5.2f%% inst %5.2f%% samples  thr=%-3s %s:%s  %s" % (100.0 * n / tot, 100.0 * smp / tots, thr, f, ln, src))
