"""Print the SASS listing of an `ncu --page source --csv` export with per-instruction executed counts
(in millions of warp instructions) and stall samples; usage: sass_hot.py file.csv [min_M]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, ismp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
tot = sum(float(r[iex]) for r in rows[2:] if len(r) > iex)
print(f"total warp instructions {tot/1e6:.1f} M")
for n, r in enumerate(rows[2:]):
    if len(r) <= iex: continue
    ex = float(r[iex]) / 1e6
    if ex >= thr:
        print(f"{n:4d} {ex:8.2f}M smp={r[ismp]:>5} {r[isrc].strip()}")
