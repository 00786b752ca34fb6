"""Per-source-line cost of a kernel from an .ncu-rep captured with --import-source on (kernel built with -lineinfo):
warp instructions executed and stall samples per CUDA source line, heaviest first.
usage: python tools/ncu_lines.py rep.ncu-rep [top_n] [file-substring]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "cuda,sass", "--csv"], stdout=subprocess.PIPE, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
only = sys.argv[3] if len(sys.argv) > 3 else ""
cur = None; rows = []; total = 0; tot_smp = 0
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) < 10 or r[0] in ("Line No", "Function Name") or not r[0].strip(): continue
    try: line, inst, smp, thr = int(r[0]), int(r[7]), int(r[6]), int(r[8])
    except ValueError: continue
    total += inst; tot_smp += smp
    if only in (cur or ""): rows.append((inst, smp, thr, cur, line, r[1].strip()[:110]))
rows.sort(reverse=True)
print(f"total warp instructions {total:,}  samples {tot_smp:,}")
for inst, smp, thr, f, line, src in rows[:top]:
    print(f"{100 * inst / total:5.1f}% inst {100 * smp / max(tot_smp, 1):5.1f}% smp  lanes {thr / max(inst, 1):4.1f}  {f}:{line}  {src}")
