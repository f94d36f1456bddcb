#!/usr/bin/env python
"""filter_proposals at the cfg5 RPN shape (64 images, 5 levels x 3 anchors, top 2000 per level): wall time and
the CUDA kernel table (torch.profiler) -- which share is ours, which is library top-k / indexing."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cpu_vision_b200 import detection  # noqa: E402

dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(1238)
B = 64
napl = [200 * 272 * 3, 100 * 136 * 3, 50 * 68 * 3, 25 * 34 * 3, 13 * 17 * 3]
A = sum(napl)
ctr = torch.rand(B, A, 2, generator=gen, device=dev) * torch.tensor([1088.0, 800.0], device=dev)
wh = torch.exp(torch.rand(B, A, 2, generator=gen, device=dev) * 3.5 + 2.5)
props = torch.cat([ctr - wh / 2, ctr + wh / 2], 2)
obj = torch.randn(B * A, 1, generator=gen, device=dev)
shapes = [(800, 1088)] * B
run = lambda: detection.filter_proposals(props, obj, shapes, napl, 2000, 2000, 0.7)  # noqa: E731
for _ in range(3):
    run()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    run()
b.record()
b.synchronize()
print("filter_proposals ms", a.elapsed_time(b) / 5)
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as p:
    run()
    torch.cuda.synchronize()
print(p.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
