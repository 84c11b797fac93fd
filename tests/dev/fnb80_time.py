"""Time one FasterNet block (C = 80, 64 x 40 x 40) through the fused kernel: python tests/dev/fnb80_time.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lead_yolo_b200 import modules as M

c = int(sys.argv[1]) if len(sys.argv) > 1 else 80
hw = int(sys.argv[2]) if len(sys.argv) > 2 else 40
mod = M.BasicStage(c, 1).cuda().eval()
x = torch.randn(64, c, hw, hw, device="cuda").to(memory_format=torch.channels_last).bfloat16()
with torch.no_grad():
    for _ in range(3):
        y = mod(x)
    torch.cuda.synchronize()
    torch.cuda._sleep(int(2e7))
    e = [torch.cuda.Event(enable_timing=True) for _ in range(21)]
    e[0].record()
    for i in range(20):
        y = mod(x)
        e[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(e[i].elapsed_time(e[i + 1]) * 1e3 for i in range(20))
    print(f"C={c} {hw}x{hw}: median {ts[10]:.1f} us, min {ts[0]:.1f} us")
