"""Module parity cases shared by the GPU tests, the CPU oracle tests and oracle/make_golden.py.

name -> (class name, constructor args, oracle key, input shape NCHW, signed input?)
Class names resolve to lead_yolo_b200.modules.<name> on the product side and to the reference's
models.common / models.rfa classes in the fixture generator.
"""

CASES = {
    "conv1x1_64_128": ("Conv", (64, 128, 1, 1), "Conv1", (2, 64, 40, 40), False),
    "conv3x3_64_64": ("Conv", (64, 64, 3, 1), "Conv3", (2, 64, 40, 40), False),
    "conv3x3_128_128_w20": ("Conv", (128, 128, 3, 1), "Conv3", (4, 128, 20, 20), False),
    "stage24": ("BasicStage", (24, 1), "BasicStage", (2, 24, 80, 80), True),
    "stage40": ("BasicStage", (40, 1), "BasicStage", (2, 40, 40, 40), True),
    "stage80": ("BasicStage", (80, 1), "BasicStage", (4, 80, 40, 40), True),
    "stage160": ("BasicStage", (160, 1), "BasicStage", (8, 160, 20, 20), True),
    "merge24_40": ("PatchMerging_FasterNet", (24, 40, 2, 2), "PatchMerging", (2, 24, 64, 64), True),
    "merge80_160": ("PatchMerging_FasterNet", (80, 160, 2, 2), "PatchMerging", (4, 80, 40, 40), True),
    "sppf160": ("SPPF", (160, 160, 5), "SPPF", (8, 160, 20, 20), True),
    "coordatt64": ("CoordAtt", (64, 64, 32), "CoordAtt", (4, 64, 48, 40), False),
    "coordatt256": ("CoordAtt", (256, 256, 32), "CoordAtt", (8, 256, 20, 20), False),
    "ca_bottleneck128": ("CA_Bottleneck", (128, 128, False, 1, 1.0), "CA_Bottleneck", (4, 128, 40, 40), False),
    "c3ca_168_128": ("C3_CA", (168, 128, 1, False), "C3_CA", (2, 168, 40, 40), False),
    "c3ca_512_512": ("C3_CA", (512, 512, 1, False), "C3_CA", (8, 512, 20, 20), False),
    "c3_256_256": ("C3", (256, 256, 1, False), "C3", (4, 256, 20, 20), False),
    "se128": ("SE", (128,), "SE", (16, 128, 12, 12), False),
    "rfcbam_k1_160_256": ("RFCBAMConv", (160, 256, 1, 1), "RFCBAM", (8, 160, 20, 20), False),
    "rfcbam_k3s2_128_128": ("RFCBAMConv", (128, 128, 3, 2), "RFCBAM", (4, 128, 40, 40), False),
    "rfcbam_k3s2_256_256": ("RFCBAMConv", (256, 256, 3, 2), "RFCBAM", (8, 256, 20, 20), False),
}

# Small variants whose real-reference outputs and gradients are committed under tests/golden/ (fp16 tensors).
GOLDEN = {
    "g_conv3x3": ("Conv", (16, 32, 3, 1), "Conv3", (1, 16, 12, 12), False),
    "g_stage24": ("BasicStage", (24, 1), "BasicStage", (1, 24, 16, 16), True),
    "g_stage40": ("BasicStage", (40, 1), "BasicStage", (1, 40, 12, 12), True),
    "g_merge": ("PatchMerging_FasterNet", (24, 40, 2, 2), "PatchMerging", (1, 24, 16, 16), True),
    "g_sppf": ("SPPF", (32, 32, 5), "SPPF", (1, 32, 12, 12), True),
    "g_coordatt": ("CoordAtt", (32, 32, 32), "CoordAtt", (2, 32, 12, 10), False),
    "g_c3ca": ("C3_CA", (48, 32, 1, False), "C3_CA", (1, 48, 12, 12), False),
    "g_c3": ("C3", (32, 32, 1, False), "C3", (1, 32, 12, 12), False),
    "g_rfcbam_k1": ("RFCBAMConv", (32, 48, 1, 1), "RFCBAM", (2, 32, 10, 10), False),
    "g_rfcbam_k3s2": ("RFCBAMConv", (32, 32, 3, 2), "RFCBAM", (2, 32, 12, 12), False),
}

IN_SEED = 3
GRAD_SEED = 7
BN_SEED = 1
INIT_SEED = 0


def rfcbam_stride(args):
    return args[3] if len(args) > 3 else 1
