"""Post-processing on the device: confidence filter + class-aware NMS of the decoded predictions — the call the
reference makes after the forward (`utils.general.non_max_suppression`, utils/general.py:884-994; val.py:228,
detect.py:130), here one C-ABI call (csrc/nms.cu) with static shapes and no wall-clock time limit.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lowlevel as ll
from ._lib import LeadYoloError, check, lib


def nms_static(prediction: torch.Tensor, conf_thres=0.25, iou_thres=0.45, agnostic=False, max_det=300,
               max_candidates=4096, out=None):
    """(det, count, n_candidates): det fp32 (bs, max_det, 6) = xyxy, conf, cls (zero rows past count[i]), count and
    n_candidates int32 (bs,).  Nothing is read back: usable inside a captured CUDA graph.  `out` = the three
    preallocated tensors (a serving engine passes its static buffers)."""
    if isinstance(prediction, (list, tuple)):
        prediction = prediction[0]
    if prediction.dtype != torch.float32 or not prediction.is_contiguous():
        prediction = prediction.float().contiguous()
    bs, rows, no = prediction.shape
    dev = prediction.device
    if out is None:
        det = torch.empty((bs, max_det, 6), dtype=torch.float32, device=dev)
        count = torch.empty(bs, dtype=torch.int32, device=dev)
        ncand = torch.empty(bs, dtype=torch.int32, device=dev)
    else:
        det, count, ncand = out
    fn = lib().ly_nms_workspace_bytes
    fn.restype = C.c_long
    ws = torch.empty(fn(bs, max_candidates), dtype=torch.uint8, device=dev)
    f = lib().ly_detect_nms
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p,
                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    f.restype = C.c_int
    with ll._Prof("ly_detect_nms"):
        check(f(prediction.data_ptr(), bs, rows, no, conf_thres, iou_thres, 1 if agnostic else 0, max_det,
                max_candidates, ws.data_ptr(), det.data_ptr(), count.data_ptr(), ncand.data_ptr(),
                torch.cuda.current_stream().cuda_stream), "ly_detect_nms")
    return det, count, ncand


def non_max_suppression(prediction, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False, multi_label=False,
                        labels=(), max_det=300, nm=0, max_candidates=4096):
    """Same signature and result as the reference's `non_max_suppression` (list of (n, 6) tensors per image) for the
    configuration LEAD-YOLO uses (best class per box).  The options without a kernel raise."""
    if classes is not None or labels or nm:
        raise LeadYoloError("non_max_suppression: classes / labels / masks have no kernel in lead_yolo_b200")
    if isinstance(prediction, (list, tuple)):
        prediction = prediction[0]
    if multi_label and prediction.shape[2] - 5 > 1:
        raise LeadYoloError("non_max_suppression: multi_label has no kernel in lead_yolo_b200")
    assert 0 <= conf_thres <= 1 and 0 <= iou_thres <= 1
    det, count, ncand = nms_static(prediction, conf_thres, iou_thres, agnostic, max_det, max_candidates)
    counts, ncs = count.tolist(), ncand.tolist()
    if max(ncs) > max_candidates:
        raise LeadYoloError(f"non_max_suppression: {max(ncs)} candidates above conf_thres in one image exceed "
                            f"max_candidates={max_candidates}; raise it (<= 8192) or the threshold")
    return [det[i, :n] for i, n in enumerate(counts)]
