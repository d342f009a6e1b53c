"""profiles/make_summary.py RAW.csv LAUNCHES.csv > profiles/ncu_rNN_summary.md
Markdown summary of one `ncu --set full` capture (exported with `ncu -i X.ncu-rep --page raw --csv`) plus the
launch list of the same program (`--metrics gpu__time_duration.sum --csv`)."""
import collections, csv, sys

COLS = [("us", "gpu__time_duration.sum", 1.0), ("regs", "launch__registers_per_thread", 1.0),
        ("dram rd MB", "dram__bytes_read.sum", 1.0), ("dram wr MB", "dram__bytes_write.sum", 1.0),
        ("dram %pk", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
        ("L2 hit %", "lts__t_sector_hit_rate.pct", 1.0),
        ("warps act %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
        ("issue act %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
        ("M warp inst", "smsp__inst_executed.sum", 1e-6),
        ("thr/inst", "smsp__thread_inst_executed_per_inst_executed.ratio", 1.0),
        ("M smem wavefronts", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", 1e-6),
        ("M L2 red sectors", "lts__t_sectors_op_red.sum", 1e-6)]
STALLS = ["short_scoreboard", "long_scoreboard", "wait", "barrier", "math_pipe_throttle", "not_selected", "selected",
          "lg_throttle", "mio_throttle", "branch_resolving", "dispatch_stall", "no_instructions", "membar", "sleeping"]


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def raw(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print("| kernel | " + " | ".join(c[0] for c in COLS) + " | top stalls (pc samples) |")
    print("|---|" + "---|" * (len(COLS) + 1))
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].replace("void ", "").replace("<unnamed>::", "").replace("(anonymous namespace)::", "")
        name = name.split("(")[0][:48]
        cells = []
        for _, m, sc in COLS:
            v = num(r[idx[m]]) if m in idx else None
            if v is None:
                cells.append("-")
                continue
            u = units[idx[m]]
            if u == "byte": v /= 1e6
            elif u == "Kbyte": v /= 1e3
            elif u == "Gbyte": v *= 1e3
            elif u == "ns": v /= 1e3
            elif u == "ms": v *= 1e3
            cells.append(f"{v * sc:.1f}" if abs(v * sc) < 1000 else f"{v * sc:.0f}")
        st = []
        tot = num(r[idx["smsp__pcsamp_sample_count"]]) if "smsp__pcsamp_sample_count" in idx else None
        for s in STALLS:
            k = "smsp__pcsamp_warps_issue_stalled_" + s
            if k in idx and tot:
                v = num(r[idx[k]])
                if v:
                    st.append((v / tot, s))
        st.sort(reverse=True)
        print(f"| `{name}` | " + " | ".join(cells) + " | " + ", ".join(f"{s} {100*f:.0f}%" for f, s in st[:4]) + " |")


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = num(r[vi])
        if v is None:
            continue
        n = r[ki].replace("void ", "").replace("<unnamed>::", "").split("(")[0][:60]
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("| kernel | launches | total us | avg us | share of GPU time |\n|---|---|---|---|---|")
    for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{n}` | {c} | {t/1e3:.1f} | {t/c/1e3:.1f} | {t/tot:.3f} |")


if __name__ == "__main__":
    print("## ncu --set full, one config-A view (cold caches, serialised kernels)\n")
    raw(sys.argv[1])
    if len(sys.argv) > 2:
        print("\n## launch list, three config-A views (gpu__time_duration.sum)\n")
        launches(sys.argv[2])
