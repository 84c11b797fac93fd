"""Device-agnostic test helpers (usable without CUDA and by oracle/make_golden.py)."""
import torch


def rel_l2(a, b):
    a, b = a.detach().float(), b.detach().float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def randomise_bn(module, seed=1):
    """Default BN (gamma=1, beta=0, mean=0, var=1) is ~identity and hides bugs."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    for m in module.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            dev = m.weight.device
            m.weight.data = (torch.rand(m.weight.shape, generator=g) * 0.8 + 0.6).to(dev)
            m.bias.data = (torch.randn(m.bias.shape, generator=g) * 0.2).to(dev)
            m.running_mean.data = (torch.randn(m.bias.shape, generator=g) * 0.2).to(dev)
            m.running_var.data = (torch.rand(m.bias.shape, generator=g) * 0.8 + 0.6).to(dev)
            m.eps = 1e-3
            m.momentum = 0.03


def make_input(shape, seed, signed=False):
    """Seeded CPU fp32 tensor whose values are exactly representable in bf16."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(shape, generator=g)
    if not signed:
        x = x.abs()
    return x.bfloat16().float()
