#!/usr/bin/env python
"""Summarise ncu CSV output for profiles/: per-kernel share of one forward step from a launch list
(--metrics gpu__time_duration.sum), and DRAM traffic / utilisation per launch from a --set full raw page.

  python tools/ncu_summary.py launches <launches.csv> [anchor-kernel-substring [kernel the segment must contain]]
  python tools/ncu_summary.py raw <raw.csv>
"""
import collections
import csv
import re
import sys


def launches(path, anchor="stem_", must=None):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((row["Kernel Name"], row["Grid Size"], row["Block Size"],
                         float(row["Metric Value"].replace(",", ""))))
    rows = [r for r in rows if "spin_kernel" not in r[0]]     # bench.py's clock-sampling keep-busy kernel
    idx = [i for i, r in enumerate(rows) if anchor in r[0]]
    seg = rows[idx[-2]:idx[-1]] if len(idx) >= 2 else rows
    if must is not None:     # the last anchor-to-anchor segment that contains a kernel named `must`
        for a, b in reversed(list(zip(idx[:-1], idx[1:]))):
            if any(must in r[0] for r in rows[a:b]):
                seg = rows[a:b]
                break
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
    """Per-launch table from an `ncu --set full --csv --page raw` log: duration, DRAM bytes, DRAM / tensor-pipe / issue
    utilisation, occupancy.  Column names are matched by suffix (recent ncu prefixes some with their section)."""
    lines = [l for l in open(path) if not l.startswith("==")]
    r = [row for row in csv.reader(lines) if row]
    hdr, units, rows = r[0], r[1], r[2:]

    def find(suffix):
        for i, h in enumerate(hdr):
            if h == suffix or h.endswith("." + suffix):
                return i
        return None

    want = [("us", "gpu__time_duration.sum"), ("dram_rd_MB", "dram__bytes_read.sum"), ("dram_wr_MB", "dram__bytes_write.sum"),
            ("dram_%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            ("tensor_%", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
            ("issue_%", "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
            ("warps_%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
            ("regs", "launch__registers_per_thread")]
    cols = [(n, find(m)) for n, m in want]
    kn, gs = find("Kernel Name"), find("Grid Size")

    def val(row, i, name):
        if i is None:
            return float("nan")
        try:
            f = float(row[i].replace(",", ""))
        except ValueError:
            return float("nan")
        u = units[i].lower()
        if name == "us":
            f *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(u, 1e-3)
        if name.endswith("_MB"):
            f *= {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1e-6)
        return f

    print(f"{'kernel':34s} {'grid':>14s} " + " ".join(f"{n:>10s}" for n, _ in cols))
    tot = collections.defaultdict(float)
    for row in rows:
        name = re.sub(r"\(.*", "", row[kn]).split("::")[-1].replace("void ", "")[:34]
        vals = [val(row, i, n) for n, i in cols]
        print(f"{name:34s} {row[gs]:>14s} " + " ".join(f"{v:10.1f}" for v in vals))
        for (n, _), v in zip(cols, vals):
            if n in ("us", "dram_rd_MB", "dram_wr_MB") and v == v:
                tot[n] += v
    print(f"# totals over {len(rows)} launches: {tot['us']:.1f} us, DRAM read {tot['dram_rd_MB']:.1f} MB, "
          f"write {tot['dram_wr_MB']:.1f} MB")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], *(sys.argv[3:5]))
    else:
        raw(sys.argv[2])
