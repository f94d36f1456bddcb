#!/usr/bin/env python
"""Experiment: does zero-filling grad_input image by image, right before that image's
scatter, keep the reds L2-resident?  (development aid; the conclusion is in DESIGN.md)

Baseline  : one memset of everything + one backward launch (what the C-ABI does today).
Per image : for b in images: memset(image b, all levels); backward(RoIs of image b) -- captured
            in a CUDA graph so the 16 x 5 launches cost no host time.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)
dtype = torch.bfloat16 if "bf16" in sys.argv else torch.float32
ph = 7 if "p7" in sys.argv else 14
per_img = 2000 if "p7" in sys.argv else 512
B = 16
feats, rois, grad = bench.make_cfg3(dev, dtype, 1237, B, per_img, ph)
shapes = [(B, bench.C, h, w) for h, w in bench.SHAPES]
lv = capi.fpn_level_map(rois, 2, 5)
gf = [torch.empty_like(f) for f in feats]


def baseline():
    capi.msroi_align_backward(grad, rois, lv, bench.SCALES, shapes, ph, ph, 2, False, layout=capi.NHWC, outs=gf)


rois0 = rois.clone()
rois0[:, 0] = 0
shapes1 = [(1, bench.C, h, w) for h, w in bench.SHAPES]


def per_image(group=1):
    for b in range(0, B, group):
        sl = slice(b * per_img, (b + group) * per_img)
        r = rois[sl].clone()
        r[:, 0] -= b
        capi.msroi_align_backward(grad[sl], r, lv[sl], bench.SCALES, [(group, bench.C, h, w) for h, w in bench.SHAPES],
                                  ph, ph, 2, False, layout=capi.NHWC, outs=[g[b:b + group] for g in gf])


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / iters


baseline()
ref = [g.clone() for g in gf]
print(f"baseline (memset + 1 launch): {timed(baseline):.4f} ms")
for group in (1, 2, 4):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        per_image(group)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            per_image(group)
    torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    err = max(float((a.float() - b.float()).abs().max()) for a, b in zip(gf, ref))
    print(f"per-{group}-image graph: {timed(g.replay):.4f} ms   max |diff| vs baseline {err:.3e}")
