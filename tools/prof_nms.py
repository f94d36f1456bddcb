#!/usr/bin/env python
"""One nms and one batched_nms call on the cfg2 workload (5000 boxes x 80 classes), for an ncu launch list."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
gen = torch.Generator(device=dev).manual_seed(1236)
xy = torch.rand(n, 2, generator=gen, device=dev) * 1000
boxes = torch.cat([xy, xy + torch.rand(n, 2, generator=gen, device=dev) * 200 + 10], 1)
scores = torch.randperm(n, generator=gen, device=dev).float() / n
idxs = torch.randint(0, 80, (n,), generator=gen, device=dev)
ws = capi.NmsWorkspace(n, dev)
for _ in range(2):
    k1 = capi.nms(boxes, scores, 0.5, workspace=ws)
    k2 = capi.batched_nms(boxes, scores, idxs, 0.5, capi.BNMS_AUTO, workspace=ws)
torch.cuda.synchronize()
print("ok", k1.numel(), k2.numel())
