"""Aggregate warp-stall samples and executed instructions per CUDA source line:
ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > f.csv ; python profiles/summarise_lines.py f.csv [N]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None; last = (None, 0, ''); agg = collections.OrderedDict()
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No": hdr = r; ix = {k: j for j, k in enumerate(hdr)}; continue
    if hdr is None or len(r) != len(hdr): continue
    try:
        n = int(r[ix["# Samples"]]); ie = int(r[ix["Instructions Executed"]])
    except ValueError:
        continue
    if r[0].strip().isdigit(): last = (cur, int(r[0]), r[1].strip())
    key = last[:2]
    a = agg.setdefault(key, [0, 0, last[2]])
    a[0] += n; a[1] += ie
tot = sum(a[0] for a in agg.values()); toti = sum(a[1] for a in agg.values())
print("total samples", tot, "warp instructions", toti)
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:N]:
    print(f"{a[0]:6d} {a[0] / tot * 100:5.1f}%  inst {a[1]:10d}  {f}:{l}  {a[2][:90]}")
