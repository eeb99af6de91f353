"""Join `ncu --page source --csv --print-source sass` (per-instruction executed counts, first launch) with the line table of
`nvdisasm -g -c` for the same kernel: dynamic opcode counts per CUDA source line.
usage: ncu_join.py src_sass.csv kernel.sass [op ...]"""
import csv, re, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = None; launch = 0; counts = []; base = None
for r in rows:
    if len(r) > 5 and r[0] == "Address":
        hdr = r; launch += 1; continue
    if hdr is None or len(r) < len(hdr) or launch != 1: continue
    try: n = int(r[hdr.index("Instructions Executed")]); s = int(r[hdr.index("# Samples")])
    except ValueError: continue
    a = int(r[0], 16)
    if base is None: base = a
    counts.append((a - base, r[1].strip(), n, s))
line_of = {}; cur = None
for ln in open(sys.argv[2]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s*/\*([0-9a-f]+)\*/', ln)
    if m: line_of[int(m.group(1), 16)] = cur
agg = collections.defaultdict(collections.Counter); samp = collections.Counter()
for off, src, n, s in counts:
    m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_]+)', src)
    agg[line_of.get(off)][m.group(2) if m else '?'] += n
    samp[line_of.get(off)] += s
tot = sum(sum(c.values()) for c in agg.values()); stot = sum(samp.values())
if len(sys.argv) > 3:
    for op in sys.argv[3:]:
        print('==', op)
        for n, k in sorted(((c[op], k) for k, c in agg.items() if c[op]), key=lambda t: -t[0])[:10]:
            print(f'   {n:>10d} {100*n/tot:4.1f}%  {k}')
else:
    print(f"total {tot}")
    for k, c in sorted(agg.items(), key=lambda kv: -sum(kv[1].values()))[:60]:
        n = sum(c.values())
        print(f'{str(k):42s} {n:>10d} {100*n/tot:4.1f}% samp {100*samp[k]/stot:4.1f}%  {dict(c.most_common(5))}')
