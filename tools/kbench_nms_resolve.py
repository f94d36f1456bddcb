import os, sys, torch
sys.path.insert(0, "/root/repo")
from cpu_vision_b200 import capi
dev = torch.device("cuda", 0)
def timed(fn, iters=200):
    for _ in range(10): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); b.synchronize()
    return a.elapsed_time(b) / iters
for n in (1000, 2000, 5000, 8192, 8663, 13312):
    gen = torch.Generator(device=dev).manual_seed(1236)
    xy = torch.rand(n, 2, generator=gen, device=dev) * 1000
    boxes = torch.cat([xy, xy + torch.rand(n, 2, generator=gen, device=dev) * 200 + 10], 1)
    scores = torch.randperm(n, generator=gen, device=dev).float() / n
    idxs = torch.randint(0, 80, (n,), generator=gen, device=dev)
    ws = capi.NmsWorkspace(n, dev)
    row = f"n={n}"
    for r in ("1", "0"):
        os.environ["B200DET_NMS_RESOLVE"] = r
        t1 = timed(lambda: capi.batched_nms(boxes, scores, idxs, 0.5, capi.BNMS_AUTO, workspace=ws, sync=False))
        t2 = timed(lambda: capi.nms(boxes, scores, 0.5, workspace=ws, sync=False))
        k = capi.nms(boxes, scores, 0.5, workspace=ws).numel()
        # RPN-like: dense overlaps, IoU 0.7
        row += f" | resolve={r}: batched {t1*1e3:6.1f} us, nms {t2*1e3:6.1f} us (kept {k})"
    print(row, flush=True)
# dense RPN-like set
n = 8000
gen = torch.Generator(device=dev).manual_seed(7)
c = torch.rand(n, 2, generator=gen, device=dev) * torch.tensor([1088.0, 800.0], device=dev)
wh = torch.exp(torch.rand(n, 2, generator=gen, device=dev) * 3.5 + 2.5)
boxes = torch.cat([c - wh / 2, c + wh / 2], 1); scores = torch.rand(n, generator=gen, device=dev)
ws = capi.NmsWorkspace(n, dev)
for r in ("1", "0"):
    os.environ["B200DET_NMS_RESOLVE"] = r
    t = timed(lambda: capi.nms(boxes, scores, 0.7, workspace=ws, sync=False))
    print(f"rpn-like n={n} thr 0.7 resolve={r}: {t*1e3:.1f} us kept {capi.nms(boxes, scores, 0.7, workspace=ws).numel()}")
