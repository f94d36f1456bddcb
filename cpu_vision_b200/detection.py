"""Detection post-processing around NMS, batched over all images.

Mirrors, with the reference's argument meaning and results,

* ``RegionProposalNetwork.filter_proposals`` (torchvision/models/detection/rpn.py:231-297) and
* ``RoIHeads.postprocess_detections`` (torchvision/models/detection/roi_heads.py:668-725),

which the reference runs as a Python loop over images: per image clip -> two filters with a
``boxes[keep]`` compaction each -> ``batched_nms`` (host read-back inside) -> top-n.  Here the shapes
stay static: one ``b200det_clip_filter_boxes`` pass writes the clipped boxes and a validity byte per
box for ALL images, one ``b200det_nms_segmented`` call runs every image's class-aware NMS
concurrently (one scan CTA per image), and a single read-back of the per-image counts at the very end
sizes the output lists.  (``postprocess_detections`` compacts the survivors of the score filter before
NMS -- one more read-back -- because there the filter drops ~99 % of the instances.)  Per-level top-k (``Tensor.topk``), sigmoid / softmax and the BoxCoder
arithmetic are library element-wise ops, written in the reference's operation order.

No CPU fallback: CUDA tensors only.
"""
import math
import os
from typing import List, Sequence, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor

from . import capi


def _require_cuda(t: Tensor, who: str):
    if not t.is_cuda:
        raise RuntimeError(f"cpu_vision_b200.{who}: tensors must be CUDA tensors (no CPU fallback)")


def _gather_kept(keep: Tensor, counts: Tensor, off: Sequence[int], *tensors: Tensor) -> List[List[Tensor]]:
    """Per-image results from the segmented keep buffer: ONE read-back of the S counts, one concatenation of
    the kept index ranges and one gather per result tensor; the per-image tensors are views of those (the
    reference's loops launch an indexing kernel per image and result)."""
    c = counts.tolist()
    if not c:
        return [[] for _ in tensors]
    allk = torch.cat([keep[off[s]: off[s] + c[s]] for s in range(len(c))])
    return [list(t[allk].split(c)) for t in tensors]


def _get_top_n_idx(objectness: Tensor, num_anchors_per_level: Sequence[int], pre_nms_top_n: int) -> Tensor:
    """rpn.py:231-240: top-k per level, indices into the concatenated anchors.  fp32 logits, at most 8 levels and
    k <= 2048: one launch for all images and levels (b200det_segmented_topk; equal scores in ascending index order,
    which torch.topk leaves unspecified).  Anything else: the reference's loop over the levels."""
    if (objectness.dtype == torch.float32 and objectness.is_cuda and 1 <= len(num_anchors_per_level) <= 8
            and 1 <= pre_nms_top_n <= capi.SEGMENTED_TOPK_MAX_K and objectness.shape[0] <= 65535
            and sum(num_anchors_per_level) == objectness.shape[1] and os.environ.get("B200DET_TOPK", "1") != "0"):
        return capi.segmented_topk(objectness, num_anchors_per_level, pre_nms_top_n)
    r, offset = [], 0
    for ob in objectness.split(list(num_anchors_per_level), 1):
        num_anchors = ob.shape[1]
        _, top_n_idx = ob.topk(min(pre_nms_top_n, num_anchors), dim=1)
        r.append(top_n_idx + offset)
        offset += num_anchors
    return torch.cat(r, dim=1)


def filter_proposals(proposals: Tensor, objectness: Tensor, image_shapes: List[Tuple[int, int]],
                     num_anchors_per_level: List[int], pre_nms_top_n: int, post_nms_top_n: int,
                     nms_thresh: float, score_thresh: float = 0.0, min_size: float = 1e-3,
                     ) -> Tuple[List[Tensor], List[Tensor]]:
    """rpn.py:242-297.  proposals [B, A, 4], objectness [B*A, 1] or [B, A] (logits)."""
    _require_cuda(proposals, "filter_proposals")
    num_images = proposals.shape[0]
    device = proposals.device
    objectness = objectness.detach().reshape(num_images, -1)
    top_n_idx = _get_top_n_idx(objectness, num_anchors_per_level, pre_nms_top_n)
    per = top_n_idx.shape[1]
    off = [i * per for i in range(num_images + 1)]
    if proposals.dtype == torch.float32 and objectness.dtype == torch.float32 and len(num_anchors_per_level) <= 8:
        # ONE kernel between top-k and NMS: gather of the selected anchors, sigmoid, pyramid level, clip, filters
        boxes, flat_scores, flat_levels, valid = capi.rpn_select(proposals, objectness, top_n_idx, image_shapes,
                                                                 num_anchors_per_level, min_size, score_thresh)
    else:  # other dtypes: the same steps as library ops + the clip / filter kernel
        levels = torch.cat([torch.full((n,), idx, dtype=torch.int64, device=device)
                            for idx, n in enumerate(num_anchors_per_level)], 0)
        levels = levels.reshape(1, -1).expand_as(objectness)
        batch_idx = torch.arange(num_images, device=device)[:, None]
        scores = torch.sigmoid(objectness[batch_idx, top_n_idx])
        flat_levels = levels[batch_idx, top_n_idx].reshape(-1)
        boxes, valid = capi.clip_filter_boxes(proposals[batch_idx, top_n_idx].reshape(-1, 4), scores.reshape(-1), off,
                                              image_shapes, min_size, score_thresh, strict=False)
        flat_scores = scores.reshape(-1).contiguous()
    keep, counts = capi.nms_segmented(boxes, flat_scores, flat_levels, off, nms_thresh, valid=valid,
                                      top_n=post_nms_top_n, sync=False)
    final_boxes, final_scores = _gather_kept(keep, counts, off, boxes, flat_scores)
    return final_boxes, final_scores


def decode_boxes(rel_codes: Tensor, boxes: Tensor, weights: Sequence[float],
                 bbox_xform_clip: float = math.log(1000.0 / 16)) -> Tensor:
    """BoxCoder.decode_single (models/detection/_utils.py:183-224), same operation order."""
    boxes = boxes.to(rel_codes.dtype)
    widths = boxes[:, 2] - boxes[:, 0]
    heights = boxes[:, 3] - boxes[:, 1]
    ctr_x = boxes[:, 0] + 0.5 * widths
    ctr_y = boxes[:, 1] + 0.5 * heights
    wx, wy, ww, wh = weights
    dx = rel_codes[:, 0::4] / wx
    dy = rel_codes[:, 1::4] / wy
    dw = torch.clamp(rel_codes[:, 2::4] / ww, max=bbox_xform_clip)
    dh = torch.clamp(rel_codes[:, 3::4] / wh, max=bbox_xform_clip)
    pred_ctr_x = dx * widths[:, None] + ctr_x[:, None]
    pred_ctr_y = dy * heights[:, None] + ctr_y[:, None]
    pred_w = torch.exp(dw) * widths[:, None]
    pred_h = torch.exp(dh) * heights[:, None]
    c_to_c_h = 0.5 * pred_h
    c_to_c_w = 0.5 * pred_w
    return torch.stack((pred_ctr_x - c_to_c_w, pred_ctr_y - c_to_c_h, pred_ctr_x + c_to_c_w, pred_ctr_y + c_to_c_h),
                       dim=2).flatten(1)


def postprocess_detections(class_logits: Tensor, box_regression: Tensor, proposals: List[Tensor],
                           image_shapes: List[Tuple[int, int]], score_thresh: float = 0.05, nms_thresh: float = 0.5,
                           detections_per_img: int = 100, bbox_reg_weights: Sequence[float] = (10.0, 10.0, 5.0, 5.0),
                           ) -> Tuple[List[Tensor], List[Tensor], List[Tensor]]:
    """roi_heads.py:668-725.  class_logits [R, num_classes], box_regression [R, num_classes * 4]."""
    _require_cuda(class_logits, "postprocess_detections")
    device = class_logits.device
    num_classes = class_logits.shape[-1]
    boxes_per_image = [b.shape[0] for b in proposals]
    R = sum(boxes_per_image)
    concat = torch.cat(proposals, 0)
    off = [0]
    for n in boxes_per_image:
        off.append(off[-1] + n * (num_classes - 1))
    if class_logits.dtype == torch.float32 and box_regression.dtype == torch.float32 and concat.dtype == torch.float32:
        # ONE kernel before NMS: softmax, BoxCoder.decode, clip, score and small-box filters (a warp per proposal)
        image_of = torch.repeat_interleave(torch.arange(len(boxes_per_image), dtype=torch.int32, device=device),
                                           torch.tensor(boxes_per_image, device=device))
        boxes, scores, valid = capi.roi_heads_decode(class_logits, box_regression.reshape(R, -1), concat, image_of,
                                                     image_shapes, bbox_reg_weights, math.log(1000.0 / 16), score_thresh,
                                                     1e-2)
    else:
        pred_boxes = decode_boxes(box_regression.reshape(R, -1) if R > 0 else box_regression, concat, bbox_reg_weights)
        pred_boxes = pred_boxes.reshape(R, -1, 4)
        pred_scores = F.softmax(class_logits, -1)
        # drop the background column, one instance per (proposal, class): roi_heads.py:697-705
        boxes = pred_boxes[:, 1:].reshape(-1, 4)
        scores = pred_scores[:, 1:].reshape(-1).contiguous()
        boxes, valid = capi.clip_filter_boxes(boxes, scores, off, image_shapes, 1e-2, score_thresh, strict=True)
    # The score filter drops most of the (proposal, class) instances (1000 proposals x 90 classes per image,
    # a few hundred above 0.05) and NMS is quadratic in the set size, so unlike the RPN side -- where nearly
    # every candidate survives and the shapes stay static -- the survivors are compacted first: ONE read-back of
    # the per-image survivor counts (the reference synchronises per image at boxes[inds]).
    csum = torch.cat([valid.new_zeros(1, dtype=torch.int64), valid.cumsum(0, dtype=torch.int64)])
    ends = csum[torch.tensor(off, device=device)].tolist()
    off2 = [e - ends[0] for e in ends]
    vidx = valid.nonzero().squeeze(1)
    # instance r * (num_classes - 1) + (c - 1) has label c (the reference's arange(num_classes)[1:] per proposal)
    boxes, scores, labels = boxes[vidx], scores[vidx], vidx % (num_classes - 1) + 1
    keep, counts = capi.nms_segmented(boxes, scores, labels, off2, nms_thresh, top_n=detections_per_img, sync=False)
    all_boxes, all_scores, all_labels = _gather_kept(keep, counts, off2, boxes, scores, labels)
    return all_boxes, all_scores, all_labels
