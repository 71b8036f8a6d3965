"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export: executed instructions per source line."""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg, fname = [], None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        fname = r[1].split('/')[-1]
        continue
    if len(r) < 8 or r[0] in ("Line No", ""):
        continue
    try:
        agg.append((int(r[7]), int(r[6]), fname, int(r[0]), r[1].strip()[:100]))
    except ValueError:
        pass
tot = sum(a[0] for a in agg)
print("total warp instructions", tot)
for a in sorted(agg, reverse=True)[:top]:
    print(f"{a[0] / tot * 100:5.1f}% samples={a[1]:6d} {a[2]}:{a[3]} {a[4]}")
