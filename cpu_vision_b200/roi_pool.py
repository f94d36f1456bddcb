"""roi_pool / RoIPool front-end with the reference's signature (torchvision/ops/roi_pool.py:14-72).  The op is
torch.ops.torchvision.roi_pool, served on CUDA tensors by csrc/roi_pool_sm100.cu behind the dispatcher shim
(autograd through torchvision::_roi_pool_backward).  No CPU path: a non-CUDA tensor raises.
"""
from typing import List, Union

import torch
from torch import nn, Tensor
from torch.nn.modules.utils import _pair

from . import _lib
from ._utils import check_roi_boxes_shape, convert_boxes_to_roi_format


def roi_pool(input: Tensor, boxes: Union[Tensor, List[Tensor]], output_size, spatial_scale: float = 1.0) -> Tensor:
    """Same arguments and meaning as the reference's roi_pool (roi_pool.py:15-54)."""
    _lib.load()
    check_roi_boxes_shape(boxes)
    rois = boxes
    output_size = _pair(output_size)
    if not isinstance(rois, torch.Tensor):
        rois = convert_boxes_to_roi_format(rois)
    if not input.is_cuda:
        raise RuntimeError("cpu_vision_b200.roi_pool: input must be a CUDA tensor (no CPU fallback)")
    output, _ = torch.ops.torchvision.roi_pool(input, rois, spatial_scale, output_size[0], output_size[1])
    return output


class RoIPool(nn.Module):
    """See :func:`roi_pool` (reference: roi_pool.py:57-72)."""

    def __init__(self, output_size, spatial_scale: float):
        super().__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale

    def forward(self, input: Tensor, rois: Union[Tensor, List[Tensor]]) -> Tensor:
        return roi_pool(input, rois, self.output_size, self.spatial_scale)

    def __repr__(self) -> str:
        return f"{self.__class__.__name__}(output_size={self.output_size}, spatial_scale={self.spatial_scale})"


def ps_roi_pool(input: Tensor, boxes: Union[Tensor, List[Tensor]], output_size, spatial_scale: float = 1.0) -> Tensor:
    """Position-sensitive RoI pooling (R-FCN); same arguments and meaning as the reference's ps_roi_pool
    (torchvision/ops/ps_roi_pool.py:14-49).  Returns Tensor[K, C / (oh * ow), oh, ow]."""
    _lib.load()
    check_roi_boxes_shape(boxes)
    rois = boxes
    output_size = _pair(output_size)
    if not isinstance(rois, torch.Tensor):
        rois = convert_boxes_to_roi_format(rois)
    if not input.is_cuda:
        raise RuntimeError("cpu_vision_b200.ps_roi_pool: input must be a CUDA tensor (no CPU fallback)")
    output, _ = torch.ops.torchvision.ps_roi_pool(input, rois, spatial_scale, output_size[0], output_size[1])
    return output


class PSRoIPool(nn.Module):
    """See :func:`ps_roi_pool` (reference: ps_roi_pool.py:52-66)."""

    def __init__(self, output_size, spatial_scale: float):
        super().__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale

    def forward(self, input: Tensor, rois: Union[Tensor, List[Tensor]]) -> Tensor:
        return ps_roi_pool(input, rois, self.output_size, self.spatial_scale)

    def __repr__(self) -> str:
        return f"{self.__class__.__name__}(output_size={self.output_size}, spatial_scale={self.spatial_scale})"


def ps_roi_align(input: Tensor, boxes: Union[Tensor, List[Tensor]], output_size, spatial_scale: float = 1.0,
                 sampling_ratio: int = -1) -> Tensor:
    """Position-sensitive RoIAlign (Light-Head R-CNN); same arguments and meaning as the reference's ps_roi_align
    (torchvision/ops/ps_roi_align.py:14-62).  Returns Tensor[K, C / (oh * ow), oh, ow]."""
    _lib.load()
    check_roi_boxes_shape(boxes)
    rois = boxes
    output_size = _pair(output_size)
    if not isinstance(rois, torch.Tensor):
        rois = convert_boxes_to_roi_format(rois)
    if not input.is_cuda:
        raise RuntimeError("cpu_vision_b200.ps_roi_align: input must be a CUDA tensor (no CPU fallback)")
    output, _ = torch.ops.torchvision.ps_roi_align(input, rois, spatial_scale, output_size[0], output_size[1],
                                                   sampling_ratio)
    return output


class PSRoIAlign(nn.Module):
    """See :func:`ps_roi_align` (reference: ps_roi_align.py:65-90)."""

    def __init__(self, output_size, spatial_scale: float, sampling_ratio: int):
        super().__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale
        self.sampling_ratio = sampling_ratio

    def forward(self, input: Tensor, rois: Union[Tensor, List[Tensor]]) -> Tensor:
        return ps_roi_align(input, rois, self.output_size, self.spatial_scale, self.sampling_ratio)

    def __repr__(self) -> str:
        return (f"{self.__class__.__name__}(output_size={self.output_size}, spatial_scale={self.spatial_scale}, "
                f"sampling_ratio={self.sampling_ratio})")
