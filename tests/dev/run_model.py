"""Run the bs=64 640x640 inference forward a few times (profiling driver): python tests/dev/run_model.py [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from bench import sar_like_batch
from lead_yolo_b200.model import build_lead_yolo

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
model = build_lead_yolo("lead", 1, "cuda", seed=0).eval()
img = sar_like_batch(64, 640, 0).cuda()
with torch.no_grad():
    for _ in range(iters):
        out = model(img)
torch.cuda.synchronize()
print("ok", tuple(out[0].shape))
