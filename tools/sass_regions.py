"""Group the listing of tools/sass_hot.py --list into contiguous regions of equal execution count and
print each region's cost in warp-instructions per loop iteration.
usage: python tools/sass_regions.py list.txt iterations [min_cost]"""
import collections, re, sys
it = float(sys.argv[2]); min_cost = float(sys.argv[3]) if len(sys.argv) > 3 else 3.0
rows = []
for l in open(sys.argv[1]):
    m = re.match(r"\s*(\d+)\s+(\d+)\s+([\d.]+)\s+(\d+)\s+(.*)", l)
    if m: rows.append((int(m.group(1)), int(m.group(2)) / it, float(m.group(3)), int(m.group(4)), m.group(5)))
reg = []; cur = None
for k, n, t, s, txt in rows:
    if cur and abs(cur['lvl'] - n) < 0.08 * max(cur['lvl'], 0.05): cur['end'] = k; cur['ins'].append(txt); cur['smp'] += s
    else: cur = {'start': k, 'end': k, 'lvl': n, 'ins': [txt], 'smp': s}; reg.append(cur)
tot = 0
for r in reg:
    cost = r['lvl'] * len(r['ins']); tot += cost
    if cost < min_cost: continue
    c = collections.Counter(re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", i).group(2) for i in r['ins'])
    fp = sum(v for k, v in c.items() if k in ('FFMA', 'FADD', 'FMUL'))
    print(f"{r['start']:5d}-{r['end']:5d} x{r['lvl']:5.2f} n={len(r['ins']):4d} cost={cost:6.1f} fp={fp:4d} smp={r['smp']:6d} ", dict(c.most_common(8)))
print("total per iteration", round(tot, 1))
