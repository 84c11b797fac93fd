"""Dev script: per-parameter gradient agreement between (a) two identical plain-autograd runs, (b) the engine's eager
step and plain autograd, (c) the engine's graph replay and plain autograd."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lead_yolo_b200.loss import ComputeLoss
from lead_yolo_b200.model import build_lead_yolo
from lead_yolo_b200.train import HYP_DEFAULT, TrainEngine, scale_hyp

g = torch.Generator().manual_seed(9)
img = torch.randint(0, 256, (4, 3, 128, 128), generator=g, dtype=torch.uint8).cuda()
tg = torch.tensor([[i, 0, 0.3 + 0.1 * i, 0.5, 0.2, 0.25] for i in range(4)], device="cuda")


def plain():
    m = build_lead_yolo("lead", 1, "cuda", seed=1).train()
    m.hyp = scale_hyp(HYP_DEFAULT, 3, 1, 128)
    loss, _ = ComputeLoss(m)(m(img), tg)
    loss.backward()
    return {n: p.grad.detach().clone() for n, p in m.named_parameters()}, float(loss)


def engine(graph):
    m = build_lead_yolo("lead", 1, "cuda", seed=1)
    e = TrainEngine(m, 4, 128, graph=graph, device="cuda", max_norm=0.0)
    e.hp_host[0] = e.hp_host[1] = e.hp_host[2] = 0.0
    loss = float(e.step(img, tg))
    names = {id(p): n for n, p in m.named_parameters()}
    return {names[id(p)]: gv.detach().clone() for p, gv in zip(e.params, e.grad_views)}, loss


def cmp(a, b, tag):
    worst = sorted(((float((a[k] - b[k]).norm() / (b[k].norm() + 1e-12)), k) for k in a), reverse=True)
    num = sum(float((a[k] - b[k]).double().pow(2).sum()) for k in a)
    den = sum(float(b[k].double().pow(2).sum()) for k in a)
    print(f"== {tag}: total rel-L2 {(num / den) ** 0.5:.3e}; worst:")
    for e, k in worst[:8]:
        print(f"   {e:.3e}  {k}  |b|={float(b[k].norm()):.3e}")


p1, l1 = plain()
p2, l2 = plain()
print("losses", l1, l2)
cmp(p1, p2, "plain vs plain (determinism)")
e1, l3 = engine(False)
cmp(e1, p1, "engine eager vs plain")
e2, l4 = engine(True)
print("losses engine", l3, l4)
cmp(e2, p1, "engine graph vs plain")
cmp(e2, e1, "engine graph vs engine eager")

print("---- variants")
from lead_yolo_b200 import ops
e3, _ = engine(False)
cmp(e3, p1, "engine eager (arena sized by earlier runs) vs plain")
_orig = ops.arena_begin
ops.arena_begin = lambda *a, **k: None
e4, _ = engine(False)
cmp(e4, p1, "engine eager, arena disabled, vs plain")
ops.arena_begin = _orig
ops._Arena.buf = None
ops._Arena.need = 0
e5, _ = engine(False)
cmp(e5, p1, "engine eager, arena reset to tiny, vs plain")
