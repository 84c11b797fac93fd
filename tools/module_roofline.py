#!/usr/bin/env python
"""Per-module roofline table of the inference forward (configs[1]) from an ncu launch list.

  python tools/module_roofline.py <launches.txt from tools/ncu_summary.py> [HBM GB/s] [tensor TFLOP/s]

The launches of one forward are consumed in order; each module takes a fixed number of them (its kernels), carries its
ALGORITHMIC bytes at the module boundary (inputs read once + outputs written once, bf16) and its flops, and
is bounded by max(bytes / HBM rate, flops / tensor rate) with the measured peaks of MEASURED_PEAKS.json.
"""
import json
import os
import re
import sys

# (module, launches, MB at the module boundary, GFLOP) at bs = 64, 640x640
MODULES = [
    ("L0 PatchEmbed (stem)", 1, 157.3, 3.78), ("L1 BasicStage 24", 1, 157.3, 8.61),
    ("L2 PatchMerging 24->40", 1, 111.4, 3.15), ("L3 BasicStage 40", 1, 65.5, 5.98),
    ("L4 PatchMerging 40->80", 1, 49.2, 2.62), ("L5 3x BasicStage 80", 3, 32.8, 17.94),
    ("L6 PatchMerging 80->160", 1, 24.6, 2.62), ("L7 BasicStage 160", 3, 16.4, 5.98), ("L8 SPPF", 3, 16.4, 3.28),
    ("L9 RFCBAMConv k1 160->256", 4, 21.3, 2.11), ("L10 Upsample", 1, 65.5, 0.0), ("L12 C3_CA 336->256", 7, 121.2, 64.61),
    ("L13 RFCBAMConv k1 256->128", 4, 78.6, 6.76), ("L14 Upsample", 1, 131.1, 0.0), ("L16 C3_CA 168->128", 7, 242.5, 64.61),
    ("L17 RFCBAMConv k3 s2 128", 4, 131.1, 32.36), ("L19 C3_CA 256->256", 7, 104.9, 60.42),
    ("L20 RFCBAMConv k3 s2 256", 4, 65.5, 31.27), ("L22 C3_CA 512->512", 7, 52.4, 60.42),
    ("L23 Detect (3 levels)", 3, 260.9, 3.30),
]


def main():
    path = sys.argv[1]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(sys.argv[2]) if len(sys.argv) > 2 else float(peaks.get("hbm_gbs", 6453.4))
    tc = float(sys.argv[3]) if len(sys.argv) > 3 else float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1432.8)))
    rows, on = [], False
    for line in open(path):
        if line.startswith("# every launch"):
            on = True
            continue
        if on and line.strip():
            m = re.match(r"(\S.*?)\s+grid .*?\s([\d.]+) us\s*$", line)
            if m:
                rows.append((m.group(1).strip(), float(m.group(2))))
    need = sum(m[1] for m in MODULES)
    assert len(rows) >= need, (len(rows), need)
    print(f"# Per-module roofline table, inference bs=64 640x640 (configs[1]).  time = sum of the module's launches in {os.path.basename(path)}")
    print(f"# (ncu gpu__time_duration, serialised, cold cache: compare shares with the CUDA-graph replay time of the bench line).")
    print(f"# bound = max(algorithmic bytes / {hbm:.1f} GB/s, algorithmic flops / {tc:.1f} TFLOP/s) at module boundaries (SURVEY §8d, MEASURED_PEAKS.json);")
    print("# Concat rows never run (K segments of the consumer GEMM).")
    print(f"{'module':30s} {'launches':>10s} {'MB':>8s} {'GFLOP':>8s} {'bound':>6s} {'bound_us':>9s} {'time_us':>8s} {'% of roof':>9s}")
    i = 0
    tb = tt = 0.0
    for name, n, mb, gf in MODULES:
        t = sum(r[1] for r in rows[i:i + n])
        i += n
        bh, bt = mb / hbm * 1e3, gf / tc * 1e3
        b = max(bh, bt)
        tb += b
        tt += t
        print(f"{name:30s} {n:10d} {mb:8.1f} {gf:8.2f} {'hbm' if bh >= bt else 'tensor':>6s} {b:9.1f} {t:8.1f} {100 * b / t:8.1f}%")
    print(f"{'model (forward + decode)':30s} {need:10d} {'':8s} {'':8s} {'':>6s} {tb:9.1f} {tt:8.1f} {100 * tb / tt:8.1f}%")
    rest = rows[need:]
    if rest:
        print("# not in the table (run only in the serving graph): " + ", ".join(f"{r[0]} {r[1]:.1f} us" for r in rest))


if __name__ == "__main__":
    main()
