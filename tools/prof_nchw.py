#!/usr/bin/env python
"""One NCHW-caller forward through the dispatcher (relayout -> fast kernel -> relayout), for ncu."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
import cpu_vision_b200  # noqa: E402,F401

dev = torch.device("cuda", 0)
feats, rois, grad = bench.make_cfg3(dev, torch.float32, 1237, 16, 512, 14)
fn = [f.contiguous() for f in feats]
del feats
out, lv = torch.ops.b200det.msroi_align(fn, rois, bench.SCALES, 14, 14, 2, False, 2, 5, 224, 4, 1e-6)
torch.cuda.synchronize()
print("ok", out.is_contiguous(), float(out.abs().mean()))
