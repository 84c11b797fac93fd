"""TEST INFRASTRUCTURE ONLY — fixtures for the loss and post-processing, made from the REAL reference
(utils/loss.py ComputeLoss, utils/general.py non_max_suppression around torchvision.ops.nms), imported unmodified
from /root/reference through oracle/ref_import.py.  Run in the build container:

    python oracle/make_golden_post.py        ->  tests/golden/loss_nms.pt
"""
from __future__ import annotations

import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_import  # noqa: E402
from oracle import restate as R  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "loss_nms.pt")
ANCHORS = torch.tensor([[10, 13, 16, 30, 33, 23], [30, 61, 62, 45, 59, 119], [116, 90, 156, 198, 373, 326]]).float()


def synthetic_targets(bs, seed, nc):
    """SSDD-shaped labels: boxes/img ~ clip(Poisson(2.2),1,29), small boxes, centres inside the image."""
    g = torch.Generator().manual_seed(seed)
    rows = []
    for b in range(bs):
        n = int(torch.poisson(torch.tensor(2.2), generator=g).clamp(1, 29))
        for _ in range(n):
            w = float(torch.normal(0.070, 0.070, (1,), generator=g).clamp(0.01, 0.5))
            h = float(torch.normal(0.116, 0.097, (1,), generator=g).clamp(0.01, 0.6))
            x = float(torch.empty(1).uniform_(w / 2, 1 - w / 2, generator=g))
            y = float(torch.empty(1).uniform_(h / 2, 1 - h / 2, generator=g))
            rows.append([b, int(torch.randint(0, nc, (1,), generator=g)), x, y, w, h])
    return torch.tensor(rows, dtype=torch.float32)


def main():
    ref = ref_import.load()
    sys.path.insert(0, ref.root)
    from utils.loss import ComputeLoss
    from utils.general import non_max_suppression
    gold = {"loss": {}, "nms": {}}
    for name, (bs, img, nc, seed) in {"nc1_128": (2, 128, 1, 0), "nc3_160": (3, 160, 3, 1), "nc1_dups": (2, 128, 1, 3)}.items():
        strides = (8, 16, 32)
        anchors = ANCHORS.view(3, 3, 2) / torch.tensor(strides).view(3, 1, 1)
        g = torch.Generator().manual_seed(100 + seed)
        p = [torch.randn((bs, 3, img // s, img // s, nc + 5), generator=g).requires_grad_(True) for s in strides]
        targets = synthetic_targets(bs, seed, nc)
        if name == "nc1_dups":   # several targets on the same cell: exercises the last-write-wins scatter
            targets = torch.cat([targets, targets[:3] + torch.tensor([0, 0, 1e-3, 1e-3, 0.01, 0.01])])
        hyp = R.scaled_hyp(3, nc, img)
        det = types.SimpleNamespace(nl=3, na=3, nc=nc, anchors=anchors, stride=torch.tensor(strides).float())
        model = types.SimpleNamespace(hyp=hyp, model=[det], parameters=lambda: iter([torch.zeros(1)]))
        loss, items = ComputeLoss(model)(p, targets)
        loss.backward()
        gold["loss"][name] = {"p": [t.detach().clone() for t in p], "targets": targets, "anchors": anchors, "hyp": hyp,
                              "nc": nc, "loss": loss.detach(), "items": items, "dp": [t.grad.clone() for t in p]}
        print(name, float(loss), items.tolist())
    for name, (bs, rows, nc, seed) in {"nc1": (3, 1008, 1, 0), "nc3": (2, 1008, 3, 1)}.items():
        g = torch.Generator().manual_seed(200 + seed)
        pred = torch.zeros(bs, rows, 5 + nc)
        # clusters of overlapping boxes around a few centres, well-spread confidences
        centres = torch.rand((bs, 12, 2), generator=g) * 500 + 60
        idx = torch.randint(0, 12, (bs, rows), generator=g)
        pred[..., 0:2] = torch.gather(centres, 1, idx[..., None].expand(-1, -1, 2)) + torch.randn((bs, rows, 2), generator=g) * 6
        pred[..., 2:4] = torch.rand((bs, rows, 2), generator=g) * 60 + 20
        pred[..., 4] = torch.rand((bs, rows), generator=g) ** 3
        pred[..., 5:] = torch.rand((bs, rows, nc), generator=g) * 0.5 + 0.5
        out = non_max_suppression(pred.clone(), 0.25, 0.45)
        gold["nms"][name] = {"pred": pred, "out": [o.clone() for o in out]}
        print(name, [o.shape[0] for o in out])
    torch.save(gold, OUT)


if __name__ == "__main__":
    main()
