#!/usr/bin/env python
"""One segmented NMS call at the cfg5 RPN shape (64 images x 8663 candidates), for an ncu launch list."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(7)
B, per = 64, int(sys.argv[1]) if len(sys.argv) > 1 else 8663
n = B * per
ctr = torch.rand(n, 2, generator=g, device=dev) * torch.tensor([1088.0, 800.0], device=dev)
wh = torch.exp(torch.rand(n, 2, generator=g, device=dev) * 3.5 + 2.5)
boxes = torch.cat([ctr - wh / 2, ctr + wh / 2], 1).clamp(min=0)
scores = torch.rand(n, generator=g, device=dev)
lv = torch.randint(0, 5, (n,), generator=g, device=dev)
off = [i * per for i in range(B + 1)]
for _ in range(2):
    keep, cnt = capi.nms_segmented(boxes, scores, lv, off, 0.7, top_n=2000, sync=False)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
keep, cnt = capi.nms_segmented(boxes, scores, lv, off, 0.7, top_n=2000, sync=False)
b.record()
b.synchronize()
print("ms", a.elapsed_time(b), "kept", cnt[:4].tolist())
