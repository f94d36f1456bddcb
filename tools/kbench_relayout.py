#!/usr/bin/env python
"""Timing of b200det_relayout on the shapes of the NCHW-caller path (cfg3 features and pooled outputs, cfg5 features),
pipelined kernels (default) against the load / barrier / store kernels (B200DET_RELAYOUT_PIPE=0).  Development aid."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters


CASES = [("feat 16x256x200x272", (16, 256, 200, 272)), ("feat 16x256x100x136", (16, 256, 100, 136)),
         ("feat 16x256x50x68", (16, 256, 50, 68)), ("feat 16x256x25x34", (16, 256, 25, 34)),
         ("pooled 8192x256x14x14", (8192, 256, 14, 14)), ("pooled 32000x256x7x7", (32000, 256, 7, 7))]
for dtype in (torch.float32, torch.bfloat16):
    for name, shape in CASES:
        x = torch.randn(shape, device=dev).to(dtype)
        cl = capi.relayout(x, True)
        assert torch.equal(cl, x)
        nbytes = 2 * x.numel() * x.element_size()
        row = f"{str(dtype).split('.')[-1]:9s} {name:24s}"
        for pipe in ("1", "0"):
            os.environ["B200DET_RELAYOUT_PIPE"] = pipe
            to_cl = timed(lambda: capi.relayout(x, True))
            to_nchw = timed(lambda: capi.relayout(cl, False))
            assert torch.equal(capi.relayout(cl, False), x)
            row += f" | pipe={pipe}: ->NHWC {to_cl:7.4f} ms {nbytes / to_cl / 1e6:6.0f} GB/s, ->NCHW {to_nchw:7.4f} ms {nbytes / to_nchw / 1e6:6.0f} GB/s"
        os.environ.pop("B200DET_RELAYOUT_PIPE", None)
        print(row, flush=True)
        del x, cl
