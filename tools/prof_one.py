#!/usr/bin/env python
"""One forward + one backward launch of a roi_align workload, for `ncu -k regex:...`.

    python tools/prof_one.py [bf16] [p7] [per_img=2000] [cfg1]
cfg3 (16 images x 512 RoIs, 14x14) by default; p7 + per_img=2000 is a 16-image slice of cfg4;
cfg1 is the single-level box-head shape (2 x 256 x 200 x 272, 1000 RoIs per image, 7x7, aligned).
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
per_img = 512
for a in sys.argv[1:]:
    if a.startswith("per_img="):
        per_img = int(a.split("=")[1])
if "cfg1" in sys.argv:
    gen = torch.Generator(device=dev).manual_seed(1235)
    x = torch.randn(2, bench.C, 200, 272, generator=gen, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last)
    r = bench.make_rois(gen, 2, 1000, 16.0, 512.0, dev)
    g = torch.randn(2000, bench.C, 7, 7, generator=gen, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last)
    out = capi.roi_align_forward(x, r, 7, 7, 0.25, 2, True)
    gi = capi.roi_align_backward(g, r, 0.25, 7, 7, 2, bench.C, 200, 272, 2, True, layout=capi.NHWC)
    torch.cuda.synchronize()
    print("ok", float(out.float().abs().mean()), float(gi.float().abs().mean()))
    sys.exit(0)
feats, rois, grad = bench.make_cfg3(dev, dtype, 1237, 16, per_img, ph)
shapes = [(16, bench.C, h, w) for h, w in bench.SHAPES]
lv = capi.fpn_level_map(rois, 2, 5)
out, _ = capi.msroi_align_forward(feats, rois, bench.SCALES, ph, ph, 2, False, levels=lv)
gf = capi.msroi_align_backward(grad, rois, lv, bench.SCALES, shapes, ph, ph, 2, False, layout=capi.NHWC)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), float(gf[0].float().abs().mean()))
