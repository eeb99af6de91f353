"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line.
usage: ncu_lines.py src.csv [topN]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = None; cur = None; agg = {}
for r in rows:
    if len(r) > 10 and r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    d = dict(zip(hdr[2:], r[2:]))
    if r[0]:
        cur = (int(r[0]), r[1].strip()[:90]); continue
    if cur is None or r[2] == "...": continue
    try:
        inst = int(d["Instructions Executed"]); samp = int(d["# Samples"])
    except ValueError:
        continue
    a = agg.setdefault(cur, [0, 0, 0, 0])
    a[0] += inst; a[1] += samp; a[2] += 1
    try: a[3] += int(d["L1 Wavefronts Shared Excessive"])
    except (ValueError, KeyError): pass
tot = sum(a[0] for a in agg.values()); tots = sum(a[1] for a in agg.values())
print(f"total warp-instructions {tot}, samples {tots}")
for (ln, src), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{ln:4d} inst {a[0]:>10d} ({100*a[0]/tot:4.1f}%) samples {100*a[1]/max(tots,1):4.1f}% sass {a[2]:3d} smemX {a[3]:>8d} | {src}")
