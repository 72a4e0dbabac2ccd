#!/usr/bin/env python3
"""Condenses ncu reports / CSV logs into the small text files kept under profiles/.
    python tools/ncu_summary.py rep <file.ncu-rep> [...]      # key metrics per captured kernel (markdown)
    python tools/ncu_summary.py launches <launches.csv> [traffic.json|-] [label]   # per-launch time, DRAM bytes, tensor %
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


def launches(path, traffic_json=None, label="fp16"):
    rows = [r for r in csv.reader(open(path)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    iname, imet, iunit, ival, iid = (hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"),
                                     hdr.index("Metric Value"), hdr.index("ID"))
    per = {}
    for r in rows[1:]:
        try:
            v = float(r[ival].replace(",", ""))
        except ValueError:
            continue
        d = per.setdefault(int(r[iid]), {"name": r[iname]})
        m = r[imet]
        if m == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(r[iunit], 1e-3)
        if m.startswith("dram__bytes"):
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[iunit], 1.0)
        d[m] = v

    def short(n):
        return n.split("(")[0].replace("void rn::<unnamed>::", "").replace("void unnamed>::", "").replace("void rn::", "")

    items = [per[k] for k in sorted(per)]
    tot = sum(d.get("gpu__time_duration.sum", 0.0) for d in items)
    print("# ncu launch list: %d launches of one forward step (batch 256, %s), %.1f us total\n" % (len(items), label, tot))
    print("Cold-cache, serialised launches: compare SHARES with the CUDA-event numbers, not absolutes.\n")
    by = {}
    for d in items:
        k = short(d["name"])
        b = by.setdefault(k, [0, 0.0, 0.0])
        b[0] += 1
        b[1] += d.get("gpu__time_duration.sum", 0.0)
        b[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    print("| kernel | launches | us | share | DRAM bytes (read+write) |\n|---|---|---|---|---|")
    for k, (c, v, t) in sorted(by.items(), key=lambda x: -x[1][1]):
        print("| %s | %d | %.1f | %.1f%% | %.1f MB |" % (k, c, v, 100 * v / tot, t / 1e6))
    print("\n| # | kernel | us | DRAM read MB | DRAM write MB | DRAM %% of peak | tensor pipe %% |\n|---|---|---|---|---|---|---|")
    for i, d in enumerate(items):
        print("| %d | %s | %.1f | %.1f | %.1f | %.1f | %.1f |" % (
            i, short(d["name"])[:60], d.get("gpu__time_duration.sum", 0.0), d.get("dram__bytes_read.sum", 0.0) / 1e6,
            d.get("dram__bytes_write.sum", 0.0) / 1e6, d.get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 0.0),
            d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0.0)))
    if traffic_json:
        import json
        conv = [d for d in items if any(t in d["name"] for t in ("conv_umma", "conv_wide", "conv_tail", "conv_pair", "halo", "stem7x7", "dualm"))]
        out = {"source": path.split("/")[-1], "conv_launches_per_step": len(conv),
               "conv_dram_bytes_per_step": sum(d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0) for d in conv),
               "all_dram_bytes_per_step": sum(d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0) for d in items),
               "conv_us_per_step_under_ncu": sum(d.get("gpu__time_duration.sum", 0.0) for d in conv)}
        json.dump(out, open(traffic_json, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "rep":
        rep(sys.argv[2:])
    else:
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 and sys.argv[3] != "-" else None, sys.argv[4] if len(sys.argv) > 4 else "fp16")
