"""The reference's OWN CUDA kernels (cuda/roi_align_kernel.cu, cuda/nms_kernel.cu, compiled unmodified for sm_100a
into baseline/_ref/tvref_cuda.so by baseline/build_ref.py --cuda) timed on this GPU, as a reported comparison
("the kernel the new implementation must beat").

Runs in a process of its own with baseline/_ref first on PYTHONPATH: the reference's package defines the schemas,
tvref_cuda.so registers the stock CUDA kernels.  The product library is never loaded here (both would register
torchvision:: CUDA kernels).  Prints one JSON object.

    PYTHONPATH=baseline/_ref python baseline/stock_cuda_bench.py
"""
import json
import os
import sys
from collections import OrderedDict

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "_ref"))
import torchvision  # noqa: E402

assert "_ref" in torchvision.__file__, torchvision.__file__
torch.ops.load_library(os.path.join(HERE, "_ref", "tvref_cuda.so"))
from torchvision import ops  # noqa: E402
from torchvision.ops.poolers import MultiScaleRoIAlign  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
SHAPES = [(200, 272), (100, 136), (50, 68), (25, 34)]


def timed(fn, iters=10, flush_l2=True):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        if flush_l2:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters


def make_rois(gen, batch, per_img, lo, hi):
    k = batch * per_img
    u = lambda: torch.rand(k, generator=gen, device=dev)
    side = torch.exp(u() * (torch.log(torch.tensor(hi / lo))) + torch.log(torch.tensor(lo)))
    aspect = torch.exp((u() - 0.5) * 1.0)
    bw, bh = (side * aspect).clamp(max=1088), (side / aspect).clamp(max=800)
    cx, cy = u() * 1088, u() * 800
    x1, y1 = (cx - bw / 2).clamp(0, 1086), (cy - bh / 2).clamp(0, 798)
    x2, y2 = torch.maximum((cx + bw / 2).clamp(max=1088), x1 + 1), torch.maximum((cy + bh / 2).clamp(max=800), y1 + 1)
    b = torch.arange(batch, device=dev).repeat_interleave(per_img).float()
    return torch.stack([b, x1, y1, x2, y2], 1).contiguous()


res = {"kernels": "cuda/roi_align_kernel.cu + cuda/nms_kernel.cu of the reference, unmodified, sm_100a",
       "gpu": torch.cuda.get_device_name(0)}
gen = torch.Generator(device=dev).manual_seed(1235)
# cfg1: 2 x 256 x 200 x 272 (NCHW, what the stock kernel takes), 1000 RoIs per image, 7x7, sr 2, aligned
x = torch.randn(2, 256, 200, 272, generator=gen, device=dev)
r = make_rois(gen, 2, 1000, 16.0, 512.0)
g = torch.randn(2000, 256, 7, 7, generator=gen, device=dev)
f = timed(lambda: torch.ops.torchvision.roi_align(x, r, 0.25, 7, 7, 2, True))
b = timed(lambda: torch.ops.torchvision._roi_align_backward(g, r, 0.25, 7, 7, 2, 256, 200, 272, 2, True))
res["cfg1_roi_align_fp32"] = {"fwd_ms": f, "bwd_ms": b, "fwd_bwd_rois_per_s": 2000 / ((f + b) / 1e3)}
del x, g
# cfg2: batched_nms 5000 boxes x 80 classes through the reference's Python (coordinate trick on CUDA) and its
# kernel (IoU mask on the device, D2H of the mask, host scan)
n = 5000
xy = torch.rand(n, 2, generator=gen, device=dev) * 1000
boxes = torch.cat([xy, xy + torch.rand(n, 2, generator=gen, device=dev) * 200 + 10], 1)
scores = torch.randperm(n, generator=gen, device=dev).float() / n
idxs = torch.randint(0, 80, (n,), generator=gen, device=dev)
ms = timed(lambda: ops.batched_nms(boxes, scores, idxs, 0.5), flush_l2=False)
res["cfg2_batched_nms_5000x80"] = {"ms": ms, "boxes_per_s": n / (ms / 1e3)}
ms = timed(lambda: ops.nms(boxes, scores, 0.5), flush_l2=False)
res["nms_5000"] = {"ms": ms, "boxes_per_s": n / (ms / 1e3)}
# cfg3 through the reference's MultiScaleRoIAlign (per-level gather -> roi_align -> index_put), NCHW features
gen = torch.Generator(device=dev).manual_seed(1237)
feats = OrderedDict((str(i), torch.randn(16, 256, h, w, generator=gen, device=dev).requires_grad_(True))
                    for i, (h, w) in enumerate(SHAPES))
rois = make_rois(gen, 16, 512, 16.0, 800.0)
blist = [rois[rois[:, 0] == i, 1:].contiguous() for i in range(16)]
grad = torch.randn(8192, 256, 14, 14, generator=gen, device=dev)
pool = MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2)
sizes = [(800, 1088)] * 16
with torch.no_grad():
    f = timed(lambda: pool(feats, blist, sizes), iters=5, flush_l2=False)


def fb():
    for t in feats.values():
        t.grad = None
    pool(feats, blist, sizes).backward(grad)


t = timed(fb, iters=5, flush_l2=False)
res["cfg3_fp32_nchw_msroi_module"] = {"fwd_ms": f, "fwd_bwd_ms": t, "fwd_bwd_rois_per_s": 8192 / (t / 1e3)}
print(json.dumps(res))
