"""Box-head layout fusion.

The reference's TwoMLPHead flattens the pooled features as NCHW, `x.flatten(start_dim=1)`
(torchvision/models/detection/faster_rcnn.py:286-307), so a channels_last roi_align output would have to be
re-laid out ([K, C, 7, 7] -> 100 MB per 2000 RoIs in fp32) before fc6.  The channels_last kernels write
[K, PH, PW, C] rows; read as they lie they are a valid fc6 input if the columns of fc6's weight are put in
the same (ph, pw, c) order once, at load time.  `flatten_pooled` is then a view -- no kernel, no copy --
and fc6 computes the same dot products (the same products summed in another order).
"""
import torch
from torch import Tensor


def flatten_pooled(x: Tensor) -> Tensor:
    """[K, C, PH, PW] -> [K, C*PH*PW] without a copy: (ph, pw, c) column order for a channels_last tensor
    (use with `fc6_weight_for_channels_last`), the reference's (c, ph, pw) order otherwise."""
    if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last) and not x.is_contiguous():
        return x.permute(0, 2, 3, 1).flatten(1)  # a view: the permuted tensor is contiguous
    return x.flatten(1)


def fc6_weight_for_channels_last(weight: Tensor, channels: int, pooled_h: int, pooled_w: int) -> Tensor:
    """fc6.weight [out, C*PH*PW] with columns in (c, ph, pw) order -> the same weights in (ph, pw, c) order."""
    out = weight.shape[0]
    if weight.shape[1] != channels * pooled_h * pooled_w:
        raise ValueError(f"fc6 weight has {weight.shape[1]} columns, expected {channels * pooled_h * pooled_w}")
    return weight.reshape(out, channels, pooled_h, pooled_w).permute(0, 2, 3, 1).reshape(out, -1).contiguous()
