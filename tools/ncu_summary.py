#!/usr/bin/env python
"""Summarise ncu CSV output for profiles/: per-kernel share of one forward step from a launch list
(--metrics gpu__time_duration.sum), and DRAM traffic / utilisation per launch from a --set full raw page.

  python tools/ncu_summary.py launches <launches.csv> [anchor-kernel-substring]
  python tools/ncu_summary.py raw <raw.csv>
"""
import collections
import csv
import re
import sys


def launches(path, anchor="stem_"):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((row["Kernel Name"], row["Grid Size"], row["Block Size"],
                         float(row["Metric Value"].replace(",", ""))))
    idx = [i for i, r in enumerate(rows) if anchor in r[0]]
    seg = rows[idx[-2]:idx[-1]] if len(idx) >= 2 else rows
    tot = sum(r[3] for r in seg)
    print(f"# one forward step (between the last two '{anchor}' launches): {len(seg)} kernels, "
          f"{tot / 1e3:.1f} us serialised, cold cache (compare SHARES, not absolutes)")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for name, _, _, ns in seg:
        short = re.sub(r"\(.*", "", name).split("::")[-1]
        agg[short][0] += 1
        agg[short][1] += ns
    print(f"{'kernel':60s} {'launches':>8s} {'us':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[:60]:60s} {v[0]:8d} {v[1] / 1e3:10.1f} {100 * v[1] / tot:6.1f}%")
    print("\n# every launch of the step, in order")
    for name, grid, block, ns in seg:
        short = re.sub(r"\(.*", "", name).split("::")[-1]
        print(f"{short[:48]:48s} grid {grid:16s} block {block:14s} {ns / 1e3:9.1f} us")


def raw(path):
    r = list(csv.reader(open(path)))
    hdr, units, rows = r[0], r[1], r[2:]
    col = {h: i for i, h in enumerate(hdr)}
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size"]
    print("kernel | " + " | ".join(f"{w} [{units[col[w]]}]" for w in want if w in col))

    def val(row, w):
        v = row[col[w]].replace(",", "")
        try:
            f = float(v)
        except ValueError:
            return 0.0
        u = units[col[w]].lower()
        if w.startswith("dram__bytes"):
            f *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        return f

    tot_r = tot_w = tot_t = 0.0
    for row in rows:
        name = re.sub(r"\(.*", "", row[col["Kernel Name"]]).split("::")[-1][:28]
        print(name + " | " + " | ".join(row[col[w]] for w in want if w in col))
        tot_r += val(row, "dram__bytes_read.sum")
        tot_w += val(row, "dram__bytes_write.sum")
        tot_t += val(row, "gpu__time_duration.sum")
    print(f"# totals over {len(rows)} launches: dram read {tot_r / 1e6:.1f} MB, write {tot_w / 1e6:.1f} MB")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], *(sys.argv[3:4]))
    else:
        raw(sys.argv[2])
