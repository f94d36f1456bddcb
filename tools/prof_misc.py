#!/usr/bin/env python
"""One launch each of the round-2 kernels outside the roi_align / nms pair, for `ncu -k regex:...`:
the pipelined re-layout (cfg3 stride-4 features NCHW -> NHWC), the per-level top-k (cfg5 RPN shape) and the push
variant of the forward (cfg4 slice, one further local buffer standing in for a peer)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(3)
x = torch.randn(16, 256, 200, 272, device=dev, generator=g)
for _ in range(2):
    cl = capi.relayout(x, True)
del x, cl
napl = [200 * 272 * 3, 100 * 136 * 3, 50 * 68 * 3, 25 * 34 * 3, 13 * 17 * 3]
obj = torch.randn(64, sum(napl), device=dev, generator=g)
for _ in range(2):
    idx = capi.segmented_topk(obj, napl, 2000)
del obj
feats, rois, _ = bench.make_cfg3(dev, torch.float32, 1237, 16, 2000, 7)
lv = capi.fpn_level_map(rois, 2, 5)
K = rois.shape[0]
mine = torch.empty(K, bench.C, 7, 7, device=dev).contiguous(memory_format=torch.channels_last)
other = torch.empty_like(mine)
for _ in range(2):
    capi.msroi_align_forward_push(feats, rois, bench.SCALES, 7, 7, 2, mine, [other.data_ptr()], levels=lv)
torch.cuda.synchronize()
print("ok", int(idx[0, 0]), bool(torch.equal(mine, other)))
