"""ctypes binding of libleadyolo.so (the C-ABI declared in include/leadyolo.h).

The library handle is a process-global loaded lazily; nothing CUDA-related is stored on nn.Module
instances, so models stay picklable / deep-copyable (reference: utils/torch_utils.py:412, train.py:390-399).
There is no fallback: if the shared library is missing, every op raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "libleadyolo.so"
_lib = None


class LeadYoloError(RuntimeError):
    pass


class ConvDesc(C.Structure):
    _fields_ = [
        ("n_src", C.c_int),
        ("src", C.c_void_p * 4),
        ("src_c", C.c_int * 4),
        ("src_ld", C.c_long * 4),
        ("N", C.c_int), ("H", C.c_int), ("W", C.c_int),
        ("kh", C.c_int), ("kw", C.c_int), ("stride", C.c_int), ("pad", C.c_int),
        ("Cout", C.c_int),
        ("w_packed", C.c_void_p),
        ("scale", C.c_void_p),
        ("shift", C.c_void_p),
        ("act", C.c_int),
        ("residual", C.c_void_p),
        ("res_ld", C.c_long),
        ("out", C.c_void_p),
        ("out_ld", C.c_long),
        ("out_fp32", C.c_int),
        ("patch_k", C.c_int), ("patch_c", C.c_int),
    ]


def lib() -> C.CDLL:
    """Load libleadyolo.so (once). Raises if it has not been built — there is no fallback path."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise LeadYoloError(
                f"{LIB_PATH} is missing: build it with `python -m lead_yolo_b200.build` "
                "(no CPU / PyTorch fallback exists for the LEAD-YOLO kernels)")
        h = C.CDLL(str(LIB_PATH))
        h.ly_last_error.restype = C.c_char_p
        _lib = h
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().ly_last_error().decode(errors="replace")
        raise LeadYoloError(f"{what} failed ({rc}): {msg}")


def stream_ptr() -> C.c_void_p:
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t) -> C.c_void_p | None:
    return None if t is None else C.c_void_p(t.data_ptr())


class PackJob(C.Structure):
    """ly_pack_job of include/leadyolo.h."""
    _fields_ = [
        ("w", C.c_void_p), ("dst", C.c_void_p),
        ("kind", C.c_int), ("Cout", C.c_int), ("Cin", C.c_int), ("kh", C.c_int), ("kw", C.c_int), ("patch", C.c_int),
        ("c", C.c_int * 4),
        ("coff", C.c_int), ("cs", C.c_int), ("P", C.c_int), ("Kp", C.c_int),
        ("total", C.c_long), ("first_block", C.c_long),
    ]


class LossDesc(C.Structure):
    """ly_loss_desc of include/leadyolo.h."""
    _fields_ = [
        ("nl", C.c_int), ("na", C.c_int), ("nc", C.c_int), ("bs", C.c_int), ("nt", C.c_int),
        ("ny", C.c_int * 3), ("nx", C.c_int * 3),
        ("p", C.c_void_p * 3),
        ("dp", C.c_void_p * 3),
        ("sb", C.c_long * 3), ("sa", C.c_long * 3), ("sy", C.c_long * 3), ("sx", C.c_long * 3), ("so", C.c_long * 3),
        ("anchors", C.c_void_p),
        ("targets", C.c_void_p),
        ("balance", C.c_float * 3),
        ("anchor_t", C.c_float), ("box", C.c_float), ("obj", C.c_float), ("cls", C.c_float),
        ("cls_pw", C.c_float), ("obj_pw", C.c_float), ("cp", C.c_float), ("cn", C.c_float),
    ]
