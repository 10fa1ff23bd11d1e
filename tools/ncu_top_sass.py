#!/usr/bin/env python
"""Top SASS instructions of an ncu capture by stall samples (optionally of one stall reason).
usage: ncu_top_sass.py <report.ncu-rep> [column, e.g. stall_long_sb] [top]"""
import csv, subprocess, sys
rep = sys.argv[1]
col = sys.argv[2] if len(sys.argv) > 2 else "# Samples"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
body = rows[2:]
tot = sum(int(r[ix[col]] or 0) for r in body); tot_s = sum(int(r[ix["# Samples"]] or 0) for r in body)
tot_i = sum(int(r[ix["Instructions Executed"]] or 0) for r in body)
print(f"total {col} = {tot}; samples {tot_s}; instructions {tot_i}")
for k in sorted(range(len(body)), key=lambda k: -int(body[k][ix[col]] or 0))[:top]:
    r = body[k]
    print(f"#{k:5d} {col}={int(r[ix[col]] or 0):7d} samples={int(r[ix['# Samples']] or 0):7d} inst={int(r[ix['Instructions Executed']] or 0):10d}  {r[ix['Source']].strip()}")
    if k > 0:
        print(f"        prev: {body[k-1][ix['Source']].strip()}")
