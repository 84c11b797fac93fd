"""Shared helpers for the parity tests: seeded inputs, BN randomisation, oracle evaluation, error metric."""
import torch

from oracle import restate as R

TOL_BF16 = 1e-2   # north_star: per-module activations and gradients within 1e-2 (rel-L2) in bf16


from helpers_cpu import make_input, randomise_bn, rel_l2  # noqa: F401


def half_normal(shape, seed, device="cuda", signed=False):
    return make_input(shape, seed, signed).to(device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)


def oracle_params(module, prefix="m", requires_grad=False, dtype=torch.float32):
    P = {}
    for k, v in module.state_dict().items():
        t = v.detach().clone()
        if t.is_floating_point():
            t = t.to(dtype)
            if requires_grad and not any(s in k for s in ("running_", "anchors")):
                t.requires_grad_(True)
        P[f"{prefix}.{k}"] = t
    return P


ORACLE_FN = {
    "Conv1": lambda P, x, c: R.conv_block(P, "m", x, c, k=1),
    "Conv3": lambda P, x, c: R.conv_block(P, "m", x, c, k=3),
    "BasicStage": lambda P, x, c: R.basic_stage(P, "m", x, c),
    "PatchEmbed": lambda P, x, c: R.patch_embed(P, "m", x, c),
    "PatchMerging": lambda P, x, c: R.patch_merge(P, "m", x, c),
    "SPPF": lambda P, x, c: R.sppf(P, "m", x, c),
    "CoordAtt": lambda P, x, c: R.coord_att(P, "m", x, c),
    "CA_Bottleneck": lambda P, x, c: R.ca_bottleneck(P, "m", x, c),
    "C3_CA": lambda P, x, c: R.c3_ca(P, "m", x, c),
    "C3": lambda P, x, c: R.c3(P, "m", x, c),
    "SE": lambda P, x, c: R.se_gate(P, "m", x),
    "RFCBAM": lambda P, x, c: R.rfcbam(P, "m", x, c),
}
