#!/usr/bin/env python3
"""Condenses ncu reports / CSV logs into the small text files kept under profiles/.
    python tools/ncu_summary.py rep <file.ncu-rep> [...]      # key metrics per captured kernel (markdown)
    python tools/ncu_summary.py launches <launches.csv>       # per-launch device time + share of the step
"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % (active cycles)"),
    ("sm__inst_executed_pipe_tensor_op_utcmma.sum", "tcgen05.mma instructions"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1/shared % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM % of peak"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__cycles_active.avg", "SMSP active cycles"),
    ("sm__cycles_elapsed.max", "SM elapsed cycles"),
]


def rep(paths):
    for path in paths:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr, units = rows[0], rows[1]
        print("## %s\n" % path.split("/")[-1])
        for r in rows[2:]:
            name = r[hdr.index("Kernel Name")]
            print("**%s**  (launch id %s)\n" % (name.replace("rn::<unnamed>::", "")[:110], r[hdr.index("ID")]))
            print("| metric | value |\n|---|---|")
            for k, label in KEYS:
                for i, h in enumerate(hdr):
                    if h == k:
                        print("| %s (`%s`) | %s %s |" % (label, k, r[i], units[i]))
            print()


def launches(path):
    rows = [r for r in csv.reader(open(path)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    iname, ival, iid = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("ID")
    items = []
    for r in rows[1:]:
        try:
            items.append((int(r[iid]), r[iname], float(r[ival].replace(",", ""))))
        except ValueError:
            continue
    unit = rows[1][hdr.index("Metric Unit")] if len(rows) > 1 else "ns"
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
    tot = sum(v for _, _, v in items) * scale
    print("# ncu launch list: %d launches, %.1f us total (cold-cache, serialised: compare shares, not absolutes)\n" % (len(items), tot))
    by = {}
    for _, n, v in items:
        key = n.split("(")[0].replace("void rn::<unnamed>::", "").replace("void rn::", "")
        by.setdefault(key, [0, 0.0])
        by[key][0] += 1
        by[key][1] += v * scale
    print("| kernel | launches | us | share |\n|---|---|---|---|")
    for k, (c, v) in sorted(by.items(), key=lambda x: -x[1][1]):
        print("| %s | %d | %.1f | %.1f%% |" % (k, c, v, 100 * v / tot))
    print("\n| # | kernel | us |\n|---|---|---|")
    for i, (idx, n, v) in enumerate(items):
        print("| %d | %s | %.1f |" % (i, n.split("(")[0].replace("void rn::<unnamed>::", "").replace("void rn::", "")[:70], v * scale))


if __name__ == "__main__":
    if sys.argv[1] == "rep":
        rep(sys.argv[2:])
    else:
        launches(sys.argv[2])
