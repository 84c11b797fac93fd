"""Dev: time ly_rf_generate_bwd pieces (dx / wgrad) at the LEAD-YOLO layer shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lead_yolo_b200 import lowlevel as ll, ops

def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

for (N, C, H, k, s) in ((64, 128, 80, 3, 2), (64, 256, 40, 3, 2), (64, 160, 20, 1, 1), (64, 256, 40, 1, 1)):
    ho, wo = ops.rf_out_size(H, H, k, s)
    x = torch.randn(N, C, H, H, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dz = torch.randn(N, k * k * C, ho, wo, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wgk = torch.randn(k * k, k * k, C, device="cuda")
    dx = torch.empty_like(x)
    dw = torch.zeros_like(wgk)
    scr = ll.rf_wgrad_scratch(C, k, "cuda")
    a = t(lambda: ll.call("ly_rf_generate_bwd", dz, x, C, N, H, H, C, k, s, ho, wo, wgk, None, dx, C, None, None, 0))
    b = t(lambda: ll.call("ly_rf_generate_bwd", dz, x, C, N, H, H, C, k, s, ho, wo, wgk, None, None, C, dw, scr, scr.numel()))
    print(f"C={C} {H}x{H} k={k} s={s}: dx {a:7.1f} us   wgrad {b:7.1f} us   (dZ {dz.numel()*2/1e6:.0f} MB)")
    if k == 3:
        wct = torch.randn(C // 16, 1296, device="cuda").bfloat16()
        zero = torch.zeros(C // 16, 144, dtype=torch.int32, device="cuda")
        r = torch.empty_like(dz, memory_format=torch.channels_last)
        c1 = t(lambda: ll.call("ly_rf_tc_taps", dz, N, ho, wo, C, wct, zero, r))
        c2 = t(lambda: ll.call("ly_rf_col2im", r, N, H, H, C, s, ho, wo, None, dx, C))
        print(f"   tensor-core dx: taps {c1:6.1f} us + col2im {c2:6.1f} us")

print("forward / gating pieces")
for (N, C, H, k, s) in ((64, 128, 80, 3, 2), (64, 256, 40, 3, 2), (64, 160, 20, 1, 1), (64, 256, 40, 1, 1)):
    ho, wo = ops.rf_out_size(H, H, k, s)
    x = torch.randn(N, C, H, H, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wgk = torch.randn(k * k, k * k, C, device="cuda")
    g = torch.empty(N, k * k * C, ho, wo, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    a_ = torch.empty_like(g, memory_format=torch.channels_last)
    da = torch.randn_like(g).contiguous(memory_format=torch.channels_last)
    dg = torch.empty_like(g, memory_format=torch.channels_last)
    se = torch.rand(N, C, 1, 1, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wa = torch.randn(18, device="cuda")
    maps = torch.empty((N, ho * k, wo * k, 2), device="cuda")
    amax = torch.empty((N, ho * k, wo * k), dtype=torch.int16, device="cuda")
    att = torch.empty((N, ho * k, wo * k), device="cuda")
    datt = torch.empty_like(att)
    dmaps = torch.empty_like(maps)
    dwa = torch.zeros(18, device="cuda")
    parts = torch.empty((ops.RF_GATE_PARTS, N, C), device="cuda")
    gen = t(lambda: ll.call("ly_rf_generate", x, C, N, H, H, C, k, s, ho, wo, wgk, None, None, ll.ACT["none"], g))
    if k == 3:
        wc = torch.randn(C // 16, 1296, device="cuda").bfloat16()
        zero = torch.zeros(C // 16, 144, dtype=torch.int32, device="cuda")
        gtc = t(lambda: ll.call("ly_rf_tc_generate", x, C, N, H, H, C, s, ho, wo, wc, zero, g))
        print(f"   tensor-core generate {gtc:6.1f} us")
    at = t(lambda: ll.call("ly_rf_attention", g, N, ho, wo, C, k, wa, maps, amax, att))
    ga = t(lambda: ll.call("ly_rf_gate", g, N, ho, wo, C, k, se, C, att, a_))
    gb = t(lambda: ll.call("ly_rf_gate_bwd", da, g, N, ho, wo, C, k, se, C, att, dg, parts, datt))
    ab = t(lambda: ll.call("ly_rf_attention_bwd", datt, att, maps, amax, N, ho, wo, C, k, wa, dmaps, dwa, dg))
    mb = g.numel() * 2 / 1e6
    print(f"C={C} {H}x{H} k={k}: generate {gen:6.1f}  poolc+att {at:6.1f}  gate {ga:6.1f}  gate_bwd {gb:6.1f}  att_bwd+poolc_bwd {ab:6.1f} us"
          f"   (G {mb:.0f} MB = {mb / 6.45e3 * 1e3:.0f} us per pass at the HBM rate)")
