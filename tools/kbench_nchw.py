#!/usr/bin/env python
"""NCHW-contiguous (torch default layout) callers: timing of the generic strided kernels, next to the
channels_last fast path on the same data.  Development aid."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)


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


for dtype in (torch.float32, torch.bfloat16):
    feats, rois, grad = bench.make_cfg3(dev, dtype, 1237, 16, 512, 14)
    shapes = [(16, bench.C, h, w) for h, w in bench.SHAPES]
    lv = capi.fpn_level_map(rois, 2, 5)
    fn = [f.contiguous() for f in feats]  # NCHW
    gn = grad.contiguous()
    t1 = timed(lambda: capi.msroi_align_forward(fn, rois, bench.SCALES, 14, 14, 2, False, levels=lv))
    t2 = timed(lambda: capi.msroi_align_backward(gn, rois, lv, bench.SCALES, shapes, 14, 14, 2, False, layout=capi.NCHW))
    t3 = timed(lambda: capi.msroi_align_forward(feats, rois, bench.SCALES, 14, 14, 2, False, levels=lv))
    t4 = timed(lambda: capi.msroi_align_backward(grad, rois, lv, bench.SCALES, shapes, 14, 14, 2, False, layout=capi.NHWC))
    import cpu_vision_b200  # noqa: F401  (dispatcher shim)
    rl = rois
    d1 = timed(lambda: torch.ops.b200det.msroi_align(fn, rl, bench.SCALES, 14, 14, 2, False, 2, 5, 224, 4, 1e-6))
    d2 = timed(lambda: torch.ops.b200det.msroi_align_backward(gn, rl, lv, bench.SCALES, [h for h, _ in bench.SHAPES],
                                                              [w for _, w in bench.SHAPES], 16, bench.C, 14, 14, 2,
                                                              False, False))
    r1 = timed(lambda: [capi.relayout(f, True) for f in fn])
    r2 = timed(lambda: capi.relayout(grad, False))
    print(f"{dtype}: dispatcher NCHW (relayout + fast kernels) fwd {d1:.3f} ms bwd {d2:.3f} ms | "
          f"b200det_relayout feats {r1:.3f} ms, out {r2:.3f} ms", flush=True)
    t5 = timed(lambda: [f.contiguous(memory_format=torch.channels_last) for f in fn])
    t6 = timed(lambda: gn.contiguous(memory_format=torch.channels_last))
    print(f"{dtype}: NCHW fwd {t1:.3f} ms bwd {t2:.3f} ms | NHWC fwd {t3:.3f} bwd {t4:.3f} | "
          f"torch relayout feats {t5:.3f} ms, grad/out {t6:.3f} ms", flush=True)
