"""Dev: which lines of lead_yolo_b200 launch torch (non-library) kernels in one training step?
   python tests/dev/train_glue.py   (on a B200)"""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lead_yolo_b200.model import build_lead_yolo          # noqa: E402
from lead_yolo_b200.train import TrainEngine              # noqa: E402

bs = int(os.environ.get("BS", 16))
m = build_lead_yolo("lead", 1, "cuda", seed=1)
eng = TrainEngine(m, bs, 640, graph=False)
img = torch.randint(0, 256, (bs, 3, 640, 640), dtype=torch.uint8, device="cuda")
tg = torch.rand(bs * 4, 6, device="cuda")
tg[:, 0] = torch.randint(0, bs, (bs * 4,), device="cuda").float()
tg[:, 1] = 0
for _ in range(2):
    eng.step(img, tg)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity   # noqa: E402
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    eng.step(img, tg)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
for ev in prof.events():
    if ev.device_time_total <= 0 or not ev.name.startswith("aten::"):
        continue
    if any(c.device_time_total > 0 and c.name.startswith("aten::") for c in ev.cpu_children):
        continue   # count leaves only
    where = "?"
    node = ev
    while node is not None and where == "?":
        for fr in node.stack:
            if "lead_yolo_b200" in fr and "torch/" not in fr:
                where = fr.split("lead_yolo_b200/")[-1]
                break
        if where == "?" and node is not ev and ("Backward" in node.name or "Fn" in node.name):
            where = node.name
        node = node.cpu_parent
    shp = ""
    key = (ev.name, where)
    agg[key][0] += 1
    agg[key][1] += ev.device_time_total
tot = sum(v[1] for v in agg.values())
print(f"torch kernels in one step: {sum(v[0] for v in agg.values())} launches, {tot:.0f} us")
for (name, where), (cnt, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
    print(f"{us:8.1f} us {cnt:4d}x  {name:28s} {where}")
