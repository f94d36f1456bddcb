"""roi_align / RoIAlign front-end with the reference's signature
(torchvision/ops/roi_align.py:203-293).  The op itself is
torch.ops.torchvision.roi_align, served on CUDA tensors by the sm_100a kernels
behind cpu_vision_b200/csrc/torch_shim.cpp.  There is no CPU or pure-PyTorch
path in this package: a non-CUDA tensor raises.
"""
from typing import List, Union

import torch
from torch import nn, Tensor
from torch.nn.modules.utils import _pair

from . import _lib
from ._utils import check_roi_boxes_shape, convert_boxes_to_roi_format


def roi_align(
    input: Tensor,
    boxes: Union[Tensor, List[Tensor]],
    output_size,
    spatial_scale: float = 1.0,
    sampling_ratio: int = -1,
    aligned: bool = False,
) -> Tensor:
    """Same arguments and meaning as the reference's roi_align (roi_align.py:204-259).

    input: Tensor[N, C, H, W] on a CUDA device (NCHW or channels_last; the
    output keeps the input's memory format).  boxes: Tensor[K, 5] or
    List[Tensor[L, 4]].  Returns Tensor[K, C, output_size[0], output_size[1]].
    """
    _lib.load()
    check_roi_boxes_shape(boxes)
    rois = boxes
    output_size = _pair(output_size)
    if not isinstance(rois, torch.Tensor):
        rois = convert_boxes_to_roi_format(rois)
    if not input.is_cuda:
        raise RuntimeError("cpu_vision_b200.roi_align: input must be a CUDA tensor (no CPU fallback)")
    return torch.ops.torchvision.roi_align(
        input, rois, spatial_scale, output_size[0], output_size[1], sampling_ratio, aligned
    )


class RoIAlign(nn.Module):
    """See :func:`roi_align` (reference: roi_align.py:262-293)."""

    def __init__(self, output_size, spatial_scale: float, sampling_ratio: int, aligned: bool = False):
        super().__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale
        self.sampling_ratio = sampling_ratio
        self.aligned = aligned

    def forward(self, input: Tensor, rois: Union[Tensor, List[Tensor]]) -> Tensor:
        return roi_align(input, rois, self.output_size, self.spatial_scale, self.sampling_ratio, self.aligned)

    def __repr__(self) -> str:
        return (
            f"{self.__class__.__name__}("
            f"output_size={self.output_size}"
            f", spatial_scale={self.spatial_scale}"
            f", sampling_ratio={self.sampling_ratio}"
            f", aligned={self.aligned}"
            f")"
        )
