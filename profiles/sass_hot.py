"""Print the SASS listing of an `ncu --page source --csv` export with per-instruction executed counts
(in millions of warp instructions) and stall samples.
usage: sass_hot.py file.csv [min_M] [section]   (section = index of the kernel when the export holds several)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
secs, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        secs.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
sec = secs[int(sys.argv[3]) if len(sys.argv) > 3 else 0]
hdr = sec["rows"][0]
isrc, ismp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
body = [r for r in sec["rows"][1:] if len(r) > iex]
tot = sum(float(r[iex]) for r in body)
print(f"{sec['name'][:100]}\ntotal warp instructions {tot/1e6:.1f} M  ({len(secs)} kernels in file)")
for n, r in enumerate(body):
    ex = float(r[iex]) / 1e6
    if ex >= thr:
        print(f"{n:4d} {ex:8.3f}M smp={r[ismp]:>5} {r[isrc].strip()}")
