"""Dev script: repeat forward+backward of single modules on identical inputs and report run-to-run differences."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lead_yolo_b200 import modules as M

torch.manual_seed(0)
CL = torch.channels_last


def hunt(tag, mod, shape, reps=60):
    mod = mod.cuda().train()
    x0 = torch.randn(shape, device="cuda").bfloat16().contiguous(memory_format=CL)
    r = torch.randn_like(mod(x0).float()) if not isinstance(mod(x0), (list, tuple)) else None
    ref = None
    bad_out = bad_dx = bad_w = 0
    worst = 0.0
    for i in range(reps):
        x = x0.clone().requires_grad_(True)
        y = mod(x)
        (y.float() * r).sum().backward()
        cur = (y.detach().clone(), x.grad.detach().clone(), [p.grad.detach().clone() for p in mod.parameters()])
        mod.zero_grad(set_to_none=True)
        if ref is None:
            ref = cur
            continue
        bad_out += int(not torch.equal(cur[0], ref[0]))
        if not torch.equal(cur[1], ref[1]):
            bad_dx += 1
            worst = max(worst, float((cur[1].float() - ref[1].float()).norm() / ref[1].float().norm()))
        for a, b in zip(cur[2], ref[2]):
            if float((a - b).norm() / (b.norm() + 1e-20)) > 1e-5:
                bad_w += 1
                break
    print(f"{tag:34s} out!= {bad_out:3d}  dx!= {bad_dx:3d} (worst rel {worst:.1e})  wgrad>1e-5 {bad_w:3d}  of {reps - 1}", flush=True)


hunt("Conv 3x3 128->128 N4 8x8", M.Conv(128, 128, 3, 1), (4, 128, 8, 8))
hunt("Conv 3x3 64->64 N4 16x16", M.Conv(64, 64, 3, 1), (4, 64, 16, 16))
hunt("Conv 3x3 256->256 N4 4x4", M.Conv(256, 256, 3, 1), (4, 256, 4, 4))
hunt("Conv 1x1 336->128 N4 8x8", M.Conv(336, 128, 1, 1), (4, 336, 8, 8))
hunt("Conv 3x3 128->128 N8 40x40", M.Conv(128, 128, 3, 1), (8, 128, 40, 40), 20)
hunt("CA_Bottleneck 128 N4 8x8", M.CA_Bottleneck(128, 128, False, 1, 1.0), (4, 128, 8, 8))
hunt("C3_CA 336->256 N4 8x8", M.C3_CA(336, 256, 1, False), (4, 336, 8, 8))
hunt("CoordAtt 128 N4 8x8", M.CoordAtt(128, 128), (4, 128, 8, 8))
hunt("RFCBAM k1 160->256 N4 4x4", M.RFCBAMConv(160, 256, 1, 1), (4, 160, 4, 4))
hunt("RFCBAM k3s2 128 N4 16x16", M.RFCBAMConv(128, 128, 3, 2), (4, 128, 16, 16))
hunt("BasicStage 160 N4 4x4", M.BasicStage(160, 1), (4, 160, 4, 4))
hunt("SPPF 160 N4 4x4", M.SPPF(160, 160, 5), (4, 160, 4, 4))
