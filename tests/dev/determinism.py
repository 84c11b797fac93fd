"""Dev script: where do two identical training-mode forwards first differ?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lead_yolo_b200 import ops
from lead_yolo_b200.model import build_lead_yolo

g = torch.Generator().manual_seed(9)
img = torch.randint(0, 256, (4, 3, 128, 128), generator=g, dtype=torch.uint8).cuda()


def run(poison=False, clear=False):
    if clear:
        ops._cache.clear()
    if poison:
        junk = [torch.full((1 << 22,), float("nan"), device="cuda") for _ in range(64)]
        del junk
    m = build_lead_yolo("lead", 1, "cuda", seed=1).train()
    outs = {}

    def hook(mod, inp, out):
        if torch.is_tensor(out):
            outs[mod.i] = out.detach().float().clone()
    hs = [c.register_forward_hook(hook) for c in m.model]
    raw = m(img)
    for h in hs:
        h.remove()
    for i, r in enumerate(raw):
        outs[100 + i] = r.detach().float().clone()
    return outs


a = run()
for tag, kw in (("plain", {}), ("clear-cache", {"clear": True}), ("poison", {"poison": True}), ("plain2", {})):
    b = run(**kw)
    first = None
    for k in sorted(a):
        d = float((a[k] - b[k]).abs().max())
        nan = bool(torch.isnan(b[k]).any())
        if (d > 0 or nan) and first is None:
            first = (k, d, nan)
    print(tag, "first differing layer:", first)
