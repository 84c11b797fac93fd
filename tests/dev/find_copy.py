"""Which torch op launches a copy kernel inside the inference forward?  python tests/dev/find_copy.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch.profiler import profile, ProfilerActivity
from bench import sar_like_batch
from lead_yolo_b200.model import build_lead_yolo

model = build_lead_yolo("lead", 1, "cuda", seed=0).eval()
img = sar_like_batch(64, 640, 0).cuda()
with torch.no_grad():
    for _ in range(3):
        model(img)
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
        model(img)
        torch.cuda.synchronize()
for e in prof.events():
    if e.name.startswith("aten::") and e.name not in ("aten::empty", "aten::empty_strided", "aten::view", "aten::as_strided",
                                                      "aten::slice", "aten::select", "aten::narrow", "aten::permute",
                                                      "aten::empty_like", "aten::reshape", "aten::_unsafe_view"):
        print(e.name, [s for s in (e.stack or []) if "lead_yolo_b200" in s][:3])
