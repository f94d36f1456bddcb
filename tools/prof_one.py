#!/usr/bin/env python
"""One forward + one (unblocked) backward launch of the cfg3 workload, for `ncu -k regex:sm100_kernel`."""
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
feats, rois, grad = bench.make_cfg3(dev, dtype, 1237, 16, 512, ph)
shapes = [(16, bench.C, h, w) for h, w in bench.SHAPES]
lv = capi.fpn_level_map(rois, 2, 5)
out, _ = capi.msroi_align_forward(feats, rois, bench.SCALES, ph, ph, 2, False, levels=lv)
gf = capi.msroi_align_backward(grad, rois, lv, bench.SCALES, shapes, ph, ph, 2, False, layout=capi.NHWC)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), float(gf[0].float().abs().mean()))
