"""BASELINE.json configs[4]: the reference's fasterrcnn_resnet50_fpn (random init, eval) on a batch of synthetic
800x1088 images with the NEW CUDA kernels under it (baseline/_ref package + cpu_vision_b200 shim), and its
post-backbone segment -- RPN filter_proposals -> MultiScaleRoIAlign -> box head -> postprocess_detections -- timed
twice: through the reference's stock per-image Python loops and through the batched mirrors
(cpu_vision_b200.detection).  Runs in a process of its own (tests/test_dropin_gpu.py explains the co-load); prints one JSON object.

    PYTHONPATH=baseline/_ref:. python baseline/cfg5_model_bench.py [batch]
"""
import json
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(HERE, "_ref"))
import torchvision  # noqa: E402

assert "_ref" in torchvision.__file__, torchvision.__file__
import cpu_vision_b200  # noqa: E402
from cpu_vision_b200 import _lib, detection  # noqa: E402
from torchvision.models.detection import fasterrcnn_resnet50_fpn  # noqa: E402
from torchvision.models.detection.rpn import concat_box_prediction_layers  # noqa: E402

assert (_lib.defined_schemas() & 3) == 0
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = fasterrcnn_resnet50_fpn(weights=None, weights_backbone=None, num_classes=91, rpn_pre_nms_top_n_test=2000,
                                rpn_post_nms_top_n_test=2000).eval().to(dev)
imgs = [torch.rand(3, 800, 1088, device=dev) for _ in range(B)]


def timed(fn, iters=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / iters


res = {"batch": B, "model": "fasterrcnn_resnet50_fpn(weights=None, num_classes=91, rpn 2000/2000).eval()",
       "gpu": torch.cuda.get_device_name(0)}
with torch.no_grad():
    res["whole_model_ms"] = timed(lambda: model(imgs), iters=3)
    images, _ = model.transform(imgs)
    features = model.backbone(images.tensors)
    res["backbone_ms"] = timed(lambda: model.backbone(images.tensors), iters=3)
    rpn, heads = model.rpn, model.roi_heads
    feats = list(features.values())
    objectness, deltas = rpn.head(feats)
    anchors = rpn.anchor_generator(images, feats)
    num_images = len(anchors)
    napl = [o[0].shape[0] * o[0].shape[1] * o[0].shape[2] for o in objectness]
    objectness, deltas = concat_box_prediction_layers(objectness, deltas)
    proposals = rpn.box_coder.decode(deltas.detach(), anchors).view(num_images, -1, 4)
    shapes = images.image_sizes
    stock_fp = lambda: rpn.filter_proposals(proposals, objectness, shapes, napl)
    ours_fp = lambda: detection.filter_proposals(proposals, objectness, shapes, napl, rpn.pre_nms_top_n(),
                                                 rpn.post_nms_top_n(), rpn.nms_thresh, rpn.score_thresh, rpn.min_size)
    res["filter_proposals_stock_loop_ms"] = timed(stock_fp)
    res["filter_proposals_batched_ms"] = timed(ours_fp)
    props, _ = stock_fp()
    pool = lambda: heads.box_roi_pool(features, props, shapes)
    res["box_roi_pool_ms"] = timed(pool)
    res["rois"] = sum(p.shape[0] for p in props)
    # the same call when the backbone hands over channels_last features (model.to(memory_format=channels_last)):
    # no re-layout passes around the kernels
    from collections import OrderedDict
    feats_cl = OrderedDict((k, v.contiguous(memory_format=torch.channels_last)) for k, v in features.items())
    res["box_roi_pool_channels_last_features_ms"] = timed(lambda: heads.box_roi_pool(feats_cl, props, shapes))
    # the fused pooler (cpu_vision_b200.poolers.MultiScaleRoIAlign: level map + all levels in ONE launch, results in
    # RoI order, no where / gather / index_put chain) on the same inputs; same constructor arguments as the model's
    from cpu_vision_b200.poolers import MultiScaleRoIAlign as FusedPool

    fused = FusedPool(heads.box_roi_pool.featmap_names, heads.box_roi_pool.output_size, heads.box_roi_pool.sampling_ratio)
    res["box_roi_pool_fused_ms"] = timed(lambda: fused(features, props, shapes))
    res["box_roi_pool_fused_channels_last_features_ms"] = timed(lambda: fused(feats_cl, props, shapes))
    res["box_roi_pool_fused_max_abs_diff"] = float((fused(features, props, shapes) - pool()).abs().max())
    box_features = pool()
    head = lambda: heads.box_predictor(heads.box_head(box_features))
    res["box_head_predictor_ms"] = timed(head)
    logits, regression = head()
    stock_pp = lambda: heads.postprocess_detections(logits, regression, props, shapes)
    ours_pp = lambda: detection.postprocess_detections(logits, regression, props, shapes, heads.score_thresh,
                                                       heads.nms_thresh, heads.detections_per_img)
    res["postprocess_detections_stock_loop_ms"] = timed(stock_pp)
    res["postprocess_detections_batched_ms"] = timed(ours_pp)
    a, b = stock_pp(), ours_pp()
    res["detections"] = sum(x.shape[0] for x in a[0])
    res["detections_batched"] = sum(x.shape[0] for x in b[0])
# the whole model once more with the reference's classes routed through the mirrors (cpu_vision_b200.dropin)
from cpu_vision_b200 import dropin  # noqa: E402

with torch.no_grad():
    dropin.enable()
    res["whole_model_dropin_enabled_ms"] = timed(lambda: model(imgs), iters=3)
    dropin.disable()
res["images_per_s_whole_model_dropin_enabled"] = B / (res["whole_model_dropin_enabled_ms"] / 1e3)
res["post_backbone_stock_ms"] = (res["filter_proposals_stock_loop_ms"] + res["box_roi_pool_ms"] +
                                 res["box_head_predictor_ms"] + res["postprocess_detections_stock_loop_ms"])
res["post_backbone_batched_ms"] = (res["filter_proposals_batched_ms"] + res["box_roi_pool_fused_ms"] +
                                   res["box_head_predictor_ms"] + res["postprocess_detections_batched_ms"])
res["images_per_s_whole_model"] = B / (res["whole_model_ms"] / 1e3)
print(json.dumps(res))
