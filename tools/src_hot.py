"""Top source lines of an `ncu --page source --csv` export by executed warp instructions (development aid).
usage: src_hot.py file.csv tiles [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
tiles = float(sys.argv[2]); top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hi]; ix = {}
for i, h in enumerate(hdr): ix.setdefault(h, i)
src = []
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or not r[0]: continue
    try: n = int(r[ix["Instructions Executed"]])
    except ValueError: continue
    src.append((n, r))
tot = sum(n for n, _ in src)
print(f"total warp instructions {tot}  per tile {tot / tiles:.0f}")
src.sort(key=lambda x: -x[0])
cum = 0
for n, r in src[:top]:
    cum += n
    print(f"{r[0]:>5} {n / tiles:8.1f} {cum / tot:5.2f} samp={r[ix['# Samples']]:>6} | {r[1][:120]}")
