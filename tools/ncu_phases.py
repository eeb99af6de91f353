"""Per-phase instruction / stall-sample shares of level3_kernel from an `ncu --page source --csv --print-source cuda,sass` dump.
usage: ncu_phases.py src.csv path/to/level_kernel3.cuh"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
src = open(sys.argv[2]).read().split('\n')
hdr = None; cur = None; agg = {}
for r in rows:
    if len(r) > 10 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    d = dict(zip(hdr[2:], r[2:]))
    if r[0]: cur = (int(r[0]), r[1]); continue
    if cur is None or r[2] == "...": continue
    try: inst = int(d["Instructions Executed"]); samp = int(d["# Samples"])
    except ValueError: continue
    a = agg.setdefault(cur, [0, 0]); a[0] += inst; a[1] += samp
def find(s):
    for i, l in enumerate(src):
        if s in l: return i + 1
    return -1000
marks = [('prologue', find('level3_kernel(const')), ('stage', find('---- 1. stage')), ('blur', find('---- 2. blur')), ('resize', find('---- 3. bilinear')),
         ('reject', find('---- 4. FAST')), ('score', find('---- 5. exact')), ('nms', find('---- 6. cell-local')), ('end', find('// Generic bilinear'))]
ph = {}
def add(name, a):
    p = ph.setdefault(name, [0, 0]); p[0] += a[0]; p[1] += a[1]
for (ln, text), a in agg.items():
    body = ln < len(src) and text.strip()[:40] == src[ln - 1].strip()[:40]
    if body and ln >= marks[0][1] and ln < marks[-1][1]:
        for i in range(len(marks) - 1):
            if marks[i][1] <= ln < marks[i + 1][1]: add(marks[i][0], a)
    elif body and find('l3_arc_score(const') <= ln <= find('l3_arc_score(const') + 24: add('score', a)
    elif body and find('l3_reflect101(int') <= ln <= find('l3_reflect101(int') + 5: add('stage', a)
    elif body and find('l3_pack4(int') <= ln <= find('l3_pack4(int') + 3: add('blur', a)
    else: add('hdr:' + text.strip()[:48], a)
tot = sum(p[0] for p in ph.values()); ts = sum(p[1] for p in ph.values())
for k, p in sorted(ph.items(), key=lambda kv: -kv[1][0]):
    if p[0] * 200 > tot: print(f"{k:56s} inst {100*p[0]/tot:5.1f}%  samples {100*p[1]/ts:5.1f}%")
