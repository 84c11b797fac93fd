"""TEST INFRASTRUCTURE ONLY — imports the UNMODIFIED reference (qingqing-zijin/LEAD-YOLO) from
/root/reference so that its own nn.Modules can act as the oracle in this container.

/root/reference does not exist on the GPU box.  There, the only copy is the git-ignored `baseline/_ref/` that
tools/install_ref.py makes (it travels with the snapshot); the drop-in GPU tests and bench.py's reference arm use it
when present and skip / fall back to the port otherwise.  Nothing in the product path imports this file.  It is also
used by oracle/make_golden*.py (fixture generation, run in the build container) and by the CPU-only tests that check
the restatement in oracle/restate.py against the real reference classes.

Shims (harness side only, reference untouched): the reference imports `timm`
(models/common.py:1167,1170-1171,1408), `IPython` (common.py:24), `matplotlib`/`seaborn`
(utils/plots.py top level) which are not installed here.  Only `nn.Identity` is ever instantiated in
place of timm's DropPath on this path (drop_path == 0, common.py:1456,1505-1506).
"""
from __future__ import annotations

import os
import sys
import types

def _find_root() -> str:
    """$LEADYOLO_REFERENCE, else /root/reference (build container), else the git-ignored copy tools/install_ref.py
    leaves in baseline/_ref (the only one that exists on the GPU box)."""
    env = os.environ.get("LEADYOLO_REFERENCE")
    if env:
        return env
    here = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")
    for cand in ("/root/reference", here):
        if os.path.isfile(os.path.join(cand, "models", "common.py")):
            return cand
    return "/root/reference"


REF_ROOT = _find_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "models", "common.py"))


class _Permissive(types.ModuleType):
    """A module whose every attribute is a harmless callable/object."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        obj = _Dummy()
        setattr(self, name, obj)
        return obj


class _Dummy:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Dummy()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Dummy()

    def __iter__(self):
        return iter(())

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def _install_stubs():
    import torch.nn as nn

    if "timm" not in sys.modules:
        class DropPath(nn.Module):
            def __init__(self, drop_prob=0.0, *a, **k):
                super().__init__()
                self.drop_prob = drop_prob

            def forward(self, x):
                assert self.drop_prob == 0.0 or not self.training, "stub DropPath is identity only"
                return x

        class SqueezeExcite(nn.Module):
            def __init__(self, *a, **k):
                super().__init__()

            def forward(self, x):
                return x

        timm = types.ModuleType("timm")
        models = types.ModuleType("timm.models")
        layers = types.ModuleType("timm.models.layers")
        eff = types.ModuleType("timm.models._efficientnet_blocks")
        layers.DropPath = DropPath
        layers.trunc_normal_ = lambda t, *a, **k: t
        layers.to_2tuple = lambda v: (v, v)
        eff.SqueezeExcite = SqueezeExcite
        timm.models = models
        models.layers = layers
        models._efficientnet_blocks = eff
        sys.modules.update({"timm": timm, "timm.models": models, "timm.models.layers": layers,
                            "timm.models._efficientnet_blocks": eff})
    for name in ("IPython", "IPython.display", "matplotlib", "matplotlib.pyplot", "matplotlib.font_manager",
                 "seaborn", "thop"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = _Permissive(name)
    # matplotlib.use / rc etc. are swallowed by _Permissive


_loaded = None


def load():
    """Returns a namespace with the reference modules: .common, .rfa, .yolo, .torch_utils, .general."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    _install_stubs()
    os.environ.setdefault("TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD", "1")
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import logging
    logging.disable(logging.INFO)
    import models.common as common  # noqa: E402
    import models.rfa as rfa  # noqa: E402
    import models.yolo as yolo  # noqa: E402
    import utils.general as general  # noqa: E402
    import utils.torch_utils as torch_utils  # noqa: E402
    ns = types.SimpleNamespace(common=common, rfa=rfa, yolo=yolo, torch_utils=torch_utils, general=general,
                               root=REF_ROOT)
    _loaded = ns
    return ns


def build_model(cfg="models/LEAD-YOLO.yaml", seed=0):
    import torch
    ref = load()
    torch.manual_seed(seed)
    return ref.yolo.Model(os.path.join(REF_ROOT, cfg), ch=3, nc=1)
