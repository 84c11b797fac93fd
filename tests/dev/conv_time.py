"""Dev: time single convolutions (eval, fused BN+SiLU) — run once per LY_CONV_* knob setting.
   python tests/dev/conv_time.py            (on a B200)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lead_yolo_b200 import modules as M   # noqa: E402

CASES = [(64, 128, 128, 40, 40, 3), (64, 256, 256, 20, 20, 3), (64, 256, 256, 40, 40, 1), (64, 336, 256, 40, 40, 1),
         (64, 256, 256, 20, 20, 1), (64, 512, 256, 20, 20, 1), (64, 128, 128, 80, 80, 1), (64, 168, 128, 80, 80, 1)]


def main():
    torch.manual_seed(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for n, c1, c2, h, w, k in CASES:
        mod = M.Conv(c1, c2, k, 1).cuda().eval()
        x = torch.randn(n, c1, h, w, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        with torch.no_grad():
            for _ in range(3):
                mod(x)
            ts = []
            for _ in range(20):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                mod(x)
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b) * 1e3)
            ts.sort()
            flop = 2.0 * n * h * w * c1 * k * k * c2
            mb = 2.0 * n * h * w * (c1 + c2) / 1e6
            print(f"{os.environ.get('TAG', ''):8s} conv{k}x{k} {c1}->{c2} @{h}x{w} x{n}: median {ts[len(ts) // 2]:7.1f} us  min {ts[0]:7.1f} us"
                  f"  ({flop / ts[len(ts) // 2] / 1e6:6.0f} TFLOP/s, {mb / ts[len(ts) // 2] * 1e3:6.0f} GB/s)")


if __name__ == "__main__":
    main()
