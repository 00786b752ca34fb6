"""Summarise an `ncu --page source --csv` (SASS view) export: executed warp-instructions per opcode
and per contiguous hot region.  usage: python tools/sass_hot.py src.csv [--list]"""
import csv, sys, re, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, iex, ith, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
tot = 0; byop = collections.Counter(); ins = []
for r in rows[2:]:
    if len(r) <= iex: continue
    n = int(r[iex]); t = int(r[ith]); s = int(r[ismp]); txt = r[isrc].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", txt)
    op = m.group(2).split(".")[0] if m else "?"
    byop[op] += n; tot += n; ins.append((n, t, s, txt))
print("total warp-inst", tot, "samples", sum(i[2] for i in ins))
for op, n in byop.most_common(40): print(f"{op:10s} {n:12d} {100*n/tot:5.1f}%")
if "--list" in sys.argv:
    for k, (n, t, s, txt) in enumerate(ins): print(f"{k:5d} {n:10d} {t/max(n,1):5.1f} {s:6d}  {txt}")
