"""profiles/line_table.py OBJECT.o FUNCTION_SUBSTRING NCU_SASS.csv [TOP] > table.md
Per-SOURCE-LINE instruction / stall table of one kernel: `ncu --page source --print-source sass --csv` gives executed
instructions and stall samples per SASS instruction; `nvdisasm -g` of the same object gives the source line of every SASS
instruction (-lineinfo build); the two listings are in the same order, so they are zipped and summed per file:line."""
import collections, csv, os, re, subprocess, sys, tempfile


def sass_lines(obj, func_sub):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, capture_output=True)
    cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    out, cur, inside = [], ("?", 0), False
    for ln in txt.splitlines():
        if ln.startswith("//---") and ".text." in ln:
            inside = func_sub in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out.append((cur, m.group(2).strip()))
    return out


def main(obj, func_sub, ncu_csv, top=28):
    rows = [r for r in csv.reader(open(ncu_csv)) if len(r) > 6 and r[2].isdigit()]
    sl = sass_lines(obj, func_sub)
    assert len(rows) == len(sl), f"ncu lists {len(rows)} instructions, nvdisasm {len(sl)}: not the same build"
    src = {}
    agg = collections.OrderedDict()
    for r, (loc, ins) in zip(rows, sl):
        a = agg.setdefault(loc, [0, 0, 0, 0.0])
        a[0] += 1
        a[1] += int(r[5])   # warp-level instructions executed
        a[2] += int(r[2])   # stall samples (all)
        a[3] += int(r[6])   # thread-level instructions executed
    tot_i = sum(a[1] for a in agg.values())
    tot_s = sum(a[2] for a in agg.values())
    root = os.path.dirname(os.path.abspath(obj))
    def text(loc):
        try:
            return open(os.path.join(root, loc[0])).read().splitlines()[loc[1] - 1].strip()[:110]
        except Exception:
            return ""
    print(f"Total: {tot_i / 1e6:.1f} M warp instructions, {tot_s} stall samples, {len(rows)} SASS instructions.\n")
    print("| file:line | SASS instr | M warp instr executed | share | stall samples share | threads/instr | source |")
    print("|---|---|---|---|---|---|---|")
    for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"| {loc[0]}:{loc[1]} | {a[0]} | {a[1] / 1e6:.2f} | {100 * a[1] / tot_i:.1f} % | {100 * a[2] / max(tot_s, 1):.1f} % | "
              f"{a[3] / max(a[1], 1):.1f} | `{text(loc).replace('|', '/')}` |")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 28)
