"""Pinned host <-> device copy bandwidth, alone and both directions at once: python tests/dev/pcie_bw.py"""
import torch
n_in, n_out = 64 * 3 * 640 * 640, 64 * 25200 * 6 * 4
h_in = torch.empty(n_in, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n_out, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n_in, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps=20):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_event(e0); s2.wait_event(e0)
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for name, a, b in (("H2D alone", 1, 0), ("D2H alone", 0, 1), ("both", 1, 1)):
    run(a, b, 3)
    ms = run(a, b)
    print(f"{name}: {ms:.3f} ms/step  H2D {a * n_in / ms / 1e6:.1f} GB/s  D2H {b * n_out / ms / 1e6:.1f} GB/s")
