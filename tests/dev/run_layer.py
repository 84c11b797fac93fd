"""Run one module at bench size a few times (for ncu): python tests/dev/run_layer.py rf17|stage24|c3ca16|stem"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from lead_yolo_b200 import modules as M

which = sys.argv[1]
torch.manual_seed(0)
if which == "rf17":
    mod, shape = M.RFCBAMConv(128, 128, 3, 2), (64, 128, 80, 80)
elif which == "stage24":
    mod, shape = M.BasicStage(24, 1), (64, 24, 160, 160)
elif which == "c3ca16":
    mod, shape = M.C3_CA(168, 128, 1, False), (64, 168, 80, 80)
elif which == "stem":
    mod, shape = M.PatchEmbed_FasterNet(3, 24, 4, 4), (64, 3, 640, 640)
mod = mod.cuda().eval()
if which == "stem":
    x = torch.randint(0, 256, shape, dtype=torch.uint8, device="cuda")
else:
    x = torch.randn(shape, device="cuda").abs().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
with torch.no_grad():
    for _ in range(3):
        y = mod(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = mod(x)
    e1.record()
    torch.cuda.synchronize()
print(which, "ms per call", e0.elapsed_time(e1) / 5)
