"""Box-list helpers; same behaviour as the reference's torchvision/ops/_utils.py:18-38."""
from typing import List, Union

import torch
from torch import Tensor


def _cat(tensors: List[Tensor], dim: int = 0) -> Tensor:
    if len(tensors) == 1:
        return tensors[0]
    return torch.cat(tensors, dim)


def convert_boxes_to_roi_format(boxes: List[Tensor]) -> Tensor:
    """List[Tensor[L_i, 4]] -> Tensor[K, 5] with the image index in column 0 (_utils.py:18-25)."""
    concat_boxes = _cat([b for b in boxes], dim=0)
    ids = _cat([torch.full_like(b[:, :1], i) for i, b in enumerate(boxes)], dim=0)
    return torch.cat([ids, concat_boxes], dim=1)


def check_roi_boxes_shape(boxes: Union[Tensor, List[Tensor]]):
    """_utils.py:28-38: same assertions, same messages."""
    if isinstance(boxes, (list, tuple)):
        for _tensor in boxes:
            torch._assert(
                _tensor.size(1) == 4, "The shape of the tensor in the boxes list is not correct as List[Tensor[L, 4]]"
            )
    elif isinstance(boxes, torch.Tensor):
        torch._assert(boxes.size(1) == 5, "The boxes tensor shape is not correct as Tensor[K, 5]")
    else:
        torch._assert(False, "boxes is expected to be a Tensor[L, 5] or a List[Tensor[K, 4]]")
    return
