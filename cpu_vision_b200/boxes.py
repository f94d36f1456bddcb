"""nms / batched_nms front-end with the reference's signatures
(torchvision/ops/boxes.py:13-112).

`nms` is torch.ops.torchvision.nms.  `batched_nms` keeps the reference's two
strategies as *semantics* -- the coordinate trick (boxes.py:78-95) when
boxes.numel() <= 20000 on a CUDA tensor, the per-class loop (boxes.py:98-112)
above that (boxes.py:72) -- but both run as one device pass
(torch.ops.b200det.batched_nms): no torch.unique sync, no per-class launches.
"""
import torch
from torch import Tensor

from . import _lib

# modes of torch.ops.b200det.batched_nms (include/b200_detection_ops.h)
_AUTO, _TRICK, _VANILLA = -1, 0, 1


def nms(boxes: Tensor, scores: Tensor, iou_threshold: float) -> Tensor:
    """int64 indices of the kept boxes, sorted by decreasing score (boxes.py:13-41)."""
    _lib.load()
    if not boxes.is_cuda:
        raise RuntimeError("cpu_vision_b200.nms: boxes must be a CUDA tensor (no CPU fallback)")
    return torch.ops.torchvision.nms(boxes, scores, iou_threshold)


def batched_nms(boxes: Tensor, scores: Tensor, idxs: Tensor, iou_threshold: float) -> Tensor:
    """Class-aware NMS (boxes.py:44-75)."""
    _lib.load()
    if not boxes.is_cuda:
        raise RuntimeError("cpu_vision_b200.batched_nms: boxes must be a CUDA tensor (no CPU fallback)")
    return torch.ops.b200det.batched_nms(boxes, scores, idxs, iou_threshold, _AUTO)


def _batched_nms_coordinate_trick(boxes: Tensor, scores: Tensor, idxs: Tensor, iou_threshold: float) -> Tensor:
    """boxes.py:78-95 semantics (fp32 rounding of boxes + idxs * (max + 1) included)."""
    _lib.load()
    if boxes.numel() == 0:
        return torch.empty((0,), dtype=torch.int64, device=boxes.device)
    return torch.ops.b200det.batched_nms(boxes, scores, idxs, iou_threshold, _TRICK)


def _batched_nms_vanilla(boxes: Tensor, scores: Tensor, idxs: Tensor, iou_threshold: float) -> Tensor:
    """boxes.py:98-112 semantics (per-class NMS on the unshifted boxes)."""
    _lib.load()
    if boxes.numel() == 0:
        return torch.empty((0,), dtype=torch.int64, device=boxes.device)
    return torch.ops.b200det.batched_nms(boxes, scores, idxs, iou_threshold, _VANILLA)
