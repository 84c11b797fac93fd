"""GPU bring-up check of ly_conv_fwd against torch conv2d (run on a B200: python tests/dev/check_conv.py).

Each case runs in its own process so one faulting kernel cannot poison the others.
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

# name: (N, H, W, src_channels, Cout, k, stride, pad, act, residual, fp32_out, scale_shift)
CASES = {
    "1x1_64_64": (2, 16, 16, [64], 64, 1, 1, 0, "none", False, False, False),
    "1x1_24_48_relu": (3, 20, 20, [24], 48, 1, 1, 0, "relu", False, False, True),
    "1x1_336_256_silu": (2, 40, 40, [336], 256, 1, 1, 0, "silu", False, False, True),
    "1x1_512_512": (2, 20, 20, [512], 512, 1, 1, 0, "silu", False, False, True),
    "1x1_2src_res": (2, 20, 20, [48, 24], 24, 1, 1, 0, "none", True, False, False),
    "1x1_4src": (1, 20, 20, [80, 80, 80, 80], 160, 1, 1, 0, "silu", False, False, True),
    "1x1_detect18_fp32": (2, 20, 20, [128], 18, 1, 1, 0, "none", False, True, True),
    "3x3_64_64": (2, 16, 16, [64], 64, 3, 1, 1, "silu", False, False, True),
    "3x3_128_128_w20": (2, 20, 20, [128], 128, 3, 1, 1, "silu", False, False, True),
    "3x3_256_256_w40": (1, 40, 40, [256], 256, 3, 1, 1, "none", False, False, False),
    "3x3_64_64_w80": (1, 80, 80, [64], 64, 3, 1, 1, "relu", False, False, True),
    "patch2_24_40": (2, 32, 32, [24], 40, 2, 2, 0, "none", False, False, True),
    "patch2_80_160": (2, 16, 24, [80], 160, 2, 2, 0, "none", False, False, True),
    "patch4_8_24": (2, 32, 32, [8], 24, 4, 4, 0, "none", False, False, True),
    "big_1x1_bs64": (64, 40, 40, [256], 128, 1, 1, 0, "silu", False, False, True),
    "l12_cv12": (64, 40, 40, [336], 256, 1, 1, 0, "silu", False, False, True),
    "l1_pconv": (64, 160, 160, [8], 8, 3, 1, 1, "none", False, False, False),
    "l1_mlp3": (64, 160, 160, [48], 24, 1, 1, 0, "none", True, False, False),
    "l12_3x3": (64, 40, 40, [128], 128, 3, 1, 1, "silu", False, False, True),
    "l16_cv12": (64, 80, 80, [168], 128, 1, 1, 0, "silu", False, False, True),
}


def run_case(name):
    import torch
    import torch.nn.functional as F
    from lead_yolo_b200 import lowlevel as ll
    from lead_yolo_b200.packing import pack_conv_weight

    N, H, W, srcc, cout, k, s, p, act, use_res, fp32o, ss = CASES[name]
    torch.manual_seed(0)
    dev = "cuda"
    cin = sum(srcc)
    xs = [torch.randn(N, c, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
          for c in srcc]
    wgt = (torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5).to(torch.bfloat16).float()
    scale = (torch.rand(cout, device=dev) + 0.5) if ss else None
    shift = torch.randn(cout, device=dev) if ss else None
    patch = s == k and k > 1
    wp = pack_conv_weight(wgt, srcc, patch=patch)
    ho, wo = (H, W) if s == 1 else (H // s, W // s)
    res = torch.randn(N, cout, ho, wo, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last) \
        if use_res else None
    out = ll.conv_gemm(xs, wp, cout, k=k, stride=s, pad=p, scale=scale, shift=shift, act=act, residual=res,
                       out_fp32=fp32o)
    torch.cuda.synchronize()
    ref = F.conv2d(torch.cat([x.float() for x in xs], 1), wgt, stride=s, padding=p)
    if ss:
        ref = ref * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)
    ref = {"none": lambda t: t, "relu": F.relu, "silu": F.silu}[act](ref)
    if use_res:
        ref = ref + res.float()
    err = ((out.float() - ref).norm() / ref.norm()).item()
    mx = (out.float() - ref).abs().max().item()
    # timing
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        ll.conv_gemm(xs, wp, cout, k=k, stride=s, pad=p, scale=scale, shift=shift, act=act, residual=res, out=out)
    t0.record()
    for _ in range(10):
        ll.conv_gemm(xs, wp, cout, k=k, stride=s, pad=p, scale=scale, shift=shift, act=act, residual=res, out=out)
    t1.record()
    torch.cuda.synchronize()
    us = t0.elapsed_time(t1) * 100
    tol = 1e-2 if not fp32o else 1e-2
    print(json.dumps({"case": name, "rel_l2": err, "max_abs": mx, "us": us, "ok": bool(err < tol)}))
    return 0 if err < tol else 1


def main():
    if len(sys.argv) > 1 and sys.argv[1] != "--all":
        sys.exit(run_case(sys.argv[1]))
    bad = 0
    for name in CASES:
        t = time.time()
        try:
            r = subprocess.run([sys.executable, __file__, name], capture_output=True, text=True, timeout=180)
            tail = (r.stdout.strip().splitlines() or [""])[-1]
            if r.returncode != 0:
                bad += 1
                tail += " | rc=%d | %s" % (r.returncode, r.stderr.strip()[-600:])
        except subprocess.TimeoutExpired:
            bad += 1
            tail = json.dumps({"case": name, "ok": False, "error": "timeout"})
        print(tail, "(%.1fs)" % (time.time() - t), flush=True)
    print("FAILED" if bad else "ALL OK", bad)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
