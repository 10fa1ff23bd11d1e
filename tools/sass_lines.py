#!/usr/bin/env python
"""Aggregate an ncu SASS source page (csv) per CUDA source line using nvdisasm -g line markers.
usage: sass_lines.py <report.ncu-rep> <cubin> <kernel-substring> [top]"""
import csv, re, subprocess, sys, collections
rep, cubin, sub = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
sass = rows[2:]
dis = subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# locate the function
start = None
for i, l in enumerate(dis):
    if l.startswith("\t.text.") or l.startswith(".text."):
        pass
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        name = m.group(1)
        if start is None and sub in name :
            start = i
        elif start is not None and i > start:
            end = i
            break
else:
    end = len(dis)
cur = None
lines = []
for l in dis[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        # consecutive marker lines form an inline chain; the last one is the outermost frame
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
print("sass rows", len(sass), "disasm instrs", len(lines), file=sys.stderr)
agg = collections.defaultdict(lambda: [0, 0, 0])
n = min(len(sass), len(lines))
tot_i = tot_s = 0
for k in range(n):
    r = sass[k]
    ie = int(r[ix["Instructions Executed"]] or 0)
    sm = int(r[ix["# Samples"]] or 0)
    a = agg[lines[k]]
    a[0] += ie; a[1] += sm; a[2] += 1
    tot_i += ie; tot_s += sm
print(f"total inst {tot_i}  samples {tot_s}")
for key, (ie, sm, cnt) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{str(key):36s} inst={ie:>12d} ({100*ie/tot_i:5.1f}%) samples={sm:>8d} ({100*sm/tot_s:5.1f}%) sass={cnt}")
if len(sys.argv) > 5:
    # phase ranges: "name:lo-hi,name:lo-hi"; instructions from headers are attributed to the previous k12 line
    ranges = [(p.split(":")[0], *map(int, p.split(":")[1].split("-"))) for p in sys.argv[5].split(",")]
    ph = collections.defaultdict(lambda: [0, 0])
    last = None
    for k in range(n):
        f, ln = lines[k] if lines[k] else ("?", 0)
        if f.startswith("k12_"):
            last = ln
        name = "other"
        for nm, lo, hi in ranges:
            if last is not None and lo <= last <= hi:
                name = nm
        r = sass[k]
        ph[name][0] += int(r[ix["Instructions Executed"]] or 0)
        ph[name][1] += int(r[ix["# Samples"]] or 0)
    for nm, (ie, sm) in ph.items():
        print(f"phase {nm:8s} inst={ie:>12d} ({100*ie/tot_i:5.1f}%) samples={sm:>8d} ({100*sm/tot_s:5.1f}%)")
