#!/usr/bin/env python
"""Put a runnable copy of the UNMODIFIED reference next to the repository, in the git-ignored `baseline/_ref/`.

The reference (qingqing-zijin/LEAD-YOLO) is not a pip-installable package (no setup.py / pyproject), so the "install"
is a plain copy of the Python sources its model path needs — `models/`, `utils/`, `data/hyps/` — taken from
/root/reference where they lie.  Nothing under baseline/_ref is tracked by git (see .gitignore); it travels to the GPU
box with the gpurun snapshot, where it serves as
  * the `--impl reference` arm of bench.py (the reference's own CPU PyTorch path, `kind: "reference"`),
  * the eager-PyTorch-on-B200 bar next to the training numbers,
  * the drop-in tests that build `models/LEAD-YOLO.yaml` through the reference's own `Model(cfg)` with this
    library's classes installed, on the GPU (tests/test_dropin_gpu.py).
It also writes `baseline/_ref/lead_yolo_refsaved.pt`: a checkpoint pickled FROM THE REFERENCE CLASSES exactly as
train.py:387-403 does (`{'model': deepcopy(model).half(), ...}`), random init (the authors' weights are not in the
tree, `.MISSING_LARGE_BLOBS`), which the drop-in tests un-pickle into the replacement classes.

    python tools/install_ref.py [reference_root]
"""
from __future__ import annotations

import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DST = os.path.join(ROOT, "baseline", "_ref")


def _writable(top):
    for base, dirs, files in os.walk(top):
        os.chmod(base, 0o755)
        for f in files:
            os.chmod(os.path.join(base, f), 0o644)


def main(src="/root/reference"):
    if not os.path.isfile(os.path.join(src, "models", "common.py")):
        print(f"install_ref: no reference tree at {src}; nothing done")
        return 1
    os.makedirs(os.path.join(DST, "data"), exist_ok=True)
    for sub in ("models", "utils", os.path.join("data", "hyps")):
        d = os.path.join(DST, sub)
        if os.path.isdir(d):
            _writable(d)
            shutil.rmtree(d)
        shutil.copytree(os.path.join(src, sub), d, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
        _writable(d)          # the source tree is read-only; the copy must stay removable
    # a checkpoint saved from the reference's own classes
    sys.path.insert(0, ROOT)
    os.environ["LEADYOLO_REFERENCE"] = DST
    from copy import deepcopy

    import torch

    from oracle import ref_import
    model = ref_import.build_model(seed=0)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers_cpu import randomise_bn
    randomise_bn(model, 1)
    ckpt = {"epoch": 0, "best_fitness": None, "model": deepcopy(model).half(), "ema": None, "updates": 0,
            "optimizer": None, "opt": {}, "git": None, "date": "synthetic"}
    torch.save(ckpt, os.path.join(DST, "lead_yolo_refsaved.pt"))
    print("install_ref: reference copied to", DST)
    return 0


if __name__ == "__main__":
    sys.exit(main(*sys.argv[1:]))
