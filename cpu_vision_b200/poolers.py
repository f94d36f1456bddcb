"""MultiScaleRoIAlign with the reference's interface (torchvision/ops/poolers.py:230-327).

What changes underneath: the reference maps RoIs to FPN levels with a chain of
small tensor ops, then loops over the levels doing where -> gather ->
roi_align -> index_put (poolers.py:183-222).  Here the level mapping is one
tiny kernel and ALL levels are pooled by one launch that writes every RoI
straight to its original row (torch.ops.b200det.msroi_align); the backward is
the matching single scatter launch.
"""
import math
from typing import Dict, List, Optional, Tuple, Union

import torch
from torch import nn, Tensor

from . import _lib


class LevelMapper:
    """FPN paper eq. (1); same constructor and call as poolers.py:47-84."""

    def __init__(self, k_min: int, k_max: int, canonical_scale: int = 224, canonical_level: int = 4,
                 eps: float = 1e-6):
        self.k_min = k_min
        self.k_max = k_max
        self.s0 = canonical_scale
        self.lvl0 = canonical_level
        self.eps = eps

    def __call__(self, boxlists: List[Tensor]) -> Tensor:
        _lib.load()
        rois = _convert_to_roi_format(boxlists)
        lv = torch.ops.b200det.fpn_level_map(rois, self.k_min, self.k_max, self.s0, self.lvl0, self.eps)
        return lv.to(torch.int64)


def initLevelMapper(k_min: int, k_max: int, canonical_scale: int = 224, canonical_level: int = 4,
                    eps: float = 1e-6):
    return LevelMapper(k_min, k_max, canonical_scale, canonical_level, eps)


def _convert_to_roi_format(boxes: List[Tensor]) -> Tensor:
    """poolers.py:87-95: cat the boxes and prepend the image index column."""
    concat_boxes = torch.cat(boxes, dim=0)
    device, dtype = concat_boxes.device, concat_boxes.dtype
    ids = torch.cat(
        [torch.full_like(b[:, :1], i, dtype=dtype, layout=torch.strided, device=device) for i, b in enumerate(boxes)],
        dim=0,
    )
    return torch.cat([ids, concat_boxes], dim=1)


def _infer_scale(feature: Tensor, original_size: List[int]) -> float:
    """poolers.py:98-106: scale = 2 ** round(log2(feat / img)), fp32 like torch.tensor(x).log2().round()."""
    size = feature.shape[-2:]
    possible_scales: List[float] = []
    for s1, s2 in zip(size, original_size):
        approx_scale = float(s1) / float(s2)
        scale = 2 ** float(torch.tensor(approx_scale).log2().round())
        possible_scales.append(scale)
    return possible_scales[0]


def _setup_scales(features: List[Tensor], image_shapes: List[Tuple[int, int]], canonical_scale: int,
                  canonical_level: int) -> Tuple[List[float], LevelMapper]:
    """poolers.py:109-134."""
    if not image_shapes:
        raise ValueError("images list should not be empty")
    max_x = 0
    max_y = 0
    for shape in image_shapes:
        max_x = max(shape[0], max_x)
        max_y = max(shape[1], max_y)
    original_input_shape = (max_x, max_y)
    scales = [_infer_scale(feat, original_input_shape) for feat in features]
    lvl_min = -torch.log2(torch.tensor(scales[0], dtype=torch.float32)).item()
    lvl_max = -torch.log2(torch.tensor(scales[-1], dtype=torch.float32)).item()
    map_levels = initLevelMapper(int(lvl_min), int(lvl_max), canonical_scale=canonical_scale,
                                 canonical_level=canonical_level)
    return scales, map_levels


def _filter_input(x: Dict[str, Tensor], featmap_names: List[str]) -> List[Tensor]:
    """poolers.py:137-143."""
    return [v for k, v in x.items() if k in featmap_names]


def _multiscale_roi_align(x_filtered: List[Tensor], boxes: List[Tensor], output_size: List[int],
                          sampling_ratio: int, scales: Optional[List[float]],
                          mapper: Optional[LevelMapper]) -> Tensor:
    """poolers.py:146-227, fused.  Results are in the original RoI order."""
    if scales is None or mapper is None:
        raise ValueError("scales and mapper should not be None")
    _lib.load()
    rois = _convert_to_roi_format(boxes)
    feats = list(x_filtered)
    if not feats[0].is_cuda:
        raise RuntimeError("cpu_vision_b200.MultiScaleRoIAlign: features must be CUDA tensors (no CPU fallback)")
    # the reference calls roi_align without `aligned` -> aligned=False (poolers.py:174-181, :204-210)
    out, _levels = torch.ops.b200det.msroi_align(
        feats, rois, [float(s) for s in scales], int(output_size[0]), int(output_size[1]), int(sampling_ratio),
        False, mapper.k_min, mapper.k_max, mapper.s0, mapper.lvl0, float(mapper.eps),
    )
    return out


class MultiScaleRoIAlign(nn.Module):
    """Same constructor, forward and repr as poolers.py:230-327."""

    __annotations__ = {"scales": Optional[List[float]], "map_levels": Optional[LevelMapper]}

    def __init__(self, featmap_names: List[str], output_size: Union[int, Tuple[int], List[int]],
                 sampling_ratio: int, *, canonical_scale: int = 224, canonical_level: int = 4):
        super().__init__()
        if isinstance(output_size, int):
            output_size = (output_size, output_size)
        self.featmap_names = featmap_names
        self.sampling_ratio = sampling_ratio
        self.output_size = tuple(output_size)
        self.scales = None
        self.map_levels = None
        self.canonical_scale = canonical_scale
        self.canonical_level = canonical_level

    def forward(self, x: Dict[str, Tensor], boxes: List[Tensor], image_shapes: List[Tuple[int, int]]) -> Tensor:
        x_filtered = _filter_input(x, self.featmap_names)
        if self.scales is None or self.map_levels is None:
            self.scales, self.map_levels = _setup_scales(
                x_filtered, image_shapes, self.canonical_scale, self.canonical_level
            )
        return _multiscale_roi_align(x_filtered, boxes, self.output_size, self.sampling_ratio, self.scales,
                                     self.map_levels)

    def __repr__(self) -> str:
        return (
            f"{self.__class__.__name__}(featmap_names={self.featmap_names}, "
            f"output_size={self.output_size}, sampling_ratio={self.sampling_ratio})"
        )
