#!/usr/bin/env python
"""Attribute the per-instruction counters of an `ncu --page source --csv` dump to CUDA source lines.
  nvdisasm -g -c <cubin> > dis.txt ; ncu -i rep --page source --csv > src.csv
  python tools/sass_lines.py dis.txt src.csv <function-substring> [source-file]
Instructions are matched by order inside the function (both listings are in address order)."""
import csv
import re
import sys

dis, src, fn = sys.argv[1:4]
srcfile = sys.argv[4] if len(sys.argv) > 4 else None
lines, cur, infn = [], None, False
for ln in open(dis, errors="replace"):
    if ln.startswith("//---") and ".text." in ln:
        infn = fn in ln
        continue
    if not infn:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        inl = "inlined" in m.group(3)
        loc = (m.group(1).split("/")[-1], int(m.group(2)))
        if srcfile and loc[0] != srcfile.split("/")[-1] and cur is not None:
            continue  # instructions of inlined header code stay with the last line of the main file
        cur = loc
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        lines.append((cur, m.group(2).strip()))
rows = list(csv.reader(open(src)))
h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
H = rows[h]
ci, si = H.index("Instructions Executed"), H.index("# Samples")
data = rows[h + 1:]
assert len(data) == len(lines), (len(data), len(lines))
agg, tot, tots = {}, 0, 0
for (loc, ins), r in zip(lines, data):
    c, s = int(r[ci] or 0), int(r[si] or 0)
    a = agg.setdefault(loc, [0, 0, 0])
    a[0] += c; a[1] += s; a[2] += 1
    tot += c; tots += s
text = {}
if srcfile:
    for i, t in enumerate(open(srcfile, errors="replace"), 1):
        text[i] = t.rstrip()[:110]
print(f"total instructions {tot}, samples {tots}")
if len(sys.argv) > 5:  # region sums: "a-b,c-d,..."
    for rg in sys.argv[5].split(","):
        a, b = (int(x) for x in rg.split("-"))
        c = sum(v[0] for k, v in agg.items() if k and a <= k[1] <= b)
        sm = sum(v[1] for k, v in agg.items() if k and a <= k[1] <= b)
        print(f"lines {a:4d}-{b:<4d} inst {100 * c / tot:5.1f}%  samp {100 * sm / max(tots, 1):5.1f}%")
    sys.exit(0)
for loc, (c, s, k) in sorted(agg.items(), key=lambda kv: (kv[0] or ("", 0))):
    if c * 200 < tot and s * 200 < tots:
        continue
    t = text.get(loc[1], "") if loc and srcfile and loc[0] == srcfile.split("/")[-1] else ""
    print(f"{loc[0] if loc else '?':>14}:{loc[1] if loc else 0:<4} inst {100 * c / tot:5.1f}%  samp {100 * s / max(tots, 1):5.1f}%  sass {k:3d} | {t}")
