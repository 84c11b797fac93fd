"""Dev: training-mode BatchNorm forward / backward at the LEAD-YOLO layer sizes (run with and without LY_BN_UNFUSED=1)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lead_yolo_b200 import lowlevel as ll   # noqa: E402


def t(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


tag = "unfused" if os.environ.get("LY_BN_UNFUSED") else "fused"
for (n, c, h) in ((64, 256, 20), (64, 128, 40), (64, 256, 40), (64, 64, 80), (64, 128, 80), (64, 48, 160), (64, 1152, 40)):
    m = n * h * h
    z = torch.randn(n, c, h, h, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn_like(z).contiguous(memory_format=torch.channels_last)
    y = torch.empty_like(z, memory_format=torch.channels_last)
    dz = torch.empty_like(z, memory_format=torch.channels_last)
    g, b = torch.rand(c, device="cuda") + 0.5, torch.randn(c, device="cuda")
    rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    stats = torch.empty(4, c, device="cuda")
    s12 = torch.empty(2, c, device="cuda")
    parts = torch.empty(ll.reduce_blocks(m, c) * 2 * c, device="cuda")

    def fwd():
        acc = torch.zeros(5 * c, device="cuda")
        ll.call("ly_bn_stats", z, c, m, c, acc)
        ll.call("ly_bn_apply_train", z, c, m, c, acc, g, b, 1e-3, 0.03, rm, rv, stats, ll.ACT["silu"], y, c)

    def bwd():
        sync = torch.zeros(2, device="cuda")
        ll.call("ly_bn_act_bwd", dy, c, z, c, m, c, stats[2], stats[3], stats[0], stats[1], ll.ACT["silu"], 1, s12[0], s12[1],
                parts, sync, dz, c)

    print(f"{tag:8s} C={c:5d} {h:3d}x{h:<3d} ({m * c / 1e6:6.1f} M elements): fwd {t(fwd):7.1f} us (roof {m * c * 4 / 6.45e6:6.1f})"
          f"   bwd {t(bwd):7.1f} us (roof {m * c * 6 / 6.45e6:6.1f})")
