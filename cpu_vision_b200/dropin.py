"""Opt-in: route the REFERENCE'S detection classes through the batched / fused mirrors without editing it.

Importing cpu_vision_b200 next to the reference's package already serves torchvision.ops.roi_align / nms with the
sm_100a kernels.  The reference's Python around them still loops over images and pyramid
levels (rpn.py:268-297, roi_heads.py:693-723, ops/poolers.py:146-227); INTEGRATION.md section 2b shows the edits a
maintainer makes.  enable() applies the same three replacements at run time:

    RegionProposalNetwork.filter_proposals -> detection.filter_proposals   (one launch sequence for all images)
    RoIHeads.postprocess_detections        -> detection.postprocess_detections
    ops.poolers._multiscale_roi_align      -> poolers._multiscale_roi_align (level map + all levels in ONE launch)

Each replacement falls through to the reference's own code for anything the mirrors do not take (CPU tensors, TorchScript
/ ONNX tracing, training-mode RPN score options they do not know).  disable() restores the originals.  Results are
the reference's (tests/test_dropin_gpu.py::test_enable_dropin_gives_the_same_detections)."""
import torch

_saved = {}


def enable():
    """Patch the reference's torchvision in this process; returns the names that were replaced."""
    import torchvision
    from torchvision.models.detection import roi_heads as _rh
    from torchvision.models.detection import rpn as _rpn
    from torchvision.ops import poolers as _tvp

    from . import detection
    from . import poolers as _ourp

    if _saved:
        return sorted(_saved)

    ref_fp = _rpn.RegionProposalNetwork.filter_proposals
    ref_pp = _rh.RoIHeads.postprocess_detections
    ref_ms = _tvp._multiscale_roi_align

    def _plain(*tensors):
        return (all(t.is_cuda for t in tensors) and not torch.jit.is_scripting() and not torch.jit.is_tracing()
                and not torchvision._is_tracing())

    def filter_proposals(self, proposals, objectness, image_shapes, num_anchors_per_level):
        if not _plain(proposals, objectness):
            return ref_fp(self, proposals, objectness, image_shapes, num_anchors_per_level)
        return detection.filter_proposals(proposals, objectness, image_shapes, num_anchors_per_level,
                                          self.pre_nms_top_n(), self.post_nms_top_n(), self.nms_thresh,
                                          self.score_thresh, self.min_size)

    def postprocess_detections(self, class_logits, box_regression, proposals, image_shapes):
        if not _plain(class_logits, box_regression) or class_logits.dtype != torch.float32:
            return ref_pp(self, class_logits, box_regression, proposals, image_shapes)
        return detection.postprocess_detections(class_logits, box_regression, proposals, image_shapes,
                                                self.score_thresh, self.nms_thresh, self.detections_per_img,
                                                self.box_coder.weights)

    def _multiscale_roi_align(x_filtered, boxes, output_size, sampling_ratio, scales, mapper):
        if scales is None or mapper is None or not _plain(*x_filtered):
            return ref_ms(x_filtered, boxes, output_size, sampling_ratio, scales, mapper)
        return _ourp._multiscale_roi_align(x_filtered, boxes, output_size, sampling_ratio, scales, mapper)

    _saved["RegionProposalNetwork.filter_proposals"] = (_rpn.RegionProposalNetwork, "filter_proposals", ref_fp)
    _saved["RoIHeads.postprocess_detections"] = (_rh.RoIHeads, "postprocess_detections", ref_pp)
    _saved["ops.poolers._multiscale_roi_align"] = (_tvp, "_multiscale_roi_align", ref_ms)
    _rpn.RegionProposalNetwork.filter_proposals = filter_proposals
    _rh.RoIHeads.postprocess_detections = postprocess_detections
    _tvp._multiscale_roi_align = _multiscale_roi_align
    return sorted(_saved)


def disable():
    for owner, name, fn in _saved.values():
        setattr(owner, name, fn)
    _saved.clear()
