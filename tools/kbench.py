#!/usr/bin/env python
"""Kernel-level timing of the roi_align paths (development aid; bench.py is the judged benchmark).

    python tools/kbench.py [cfg3_fp32 cfg3_bf16 cfg1 cfg4 ...]
Prints one line per leg: ms, algorithmic GB/s and fraction of the measured HBM peak.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)
peak, _ = bench.measured_peak()
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters=10, flush_l2=True):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    best = 1e9
    for _ in range(iters):
        if flush_l2:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        t = a.elapsed_time(b)
        tot += t
        best = min(best, t)
    return tot / iters, best


def report(name, ms, best, nbytes):
    gbs = nbytes / (ms / 1e3) / 1e9
    print(f"{name:34s} {ms:8.4f} ms (best {best:8.4f})  {gbs:8.1f} GB/s  {gbs / peak * 100:5.1f}% of {peak:.0f}", flush=True)


def cfg3(dtype, batch=16, per_img=512, ph=14, tag="cfg3"):
    feats, rois, grad = bench.make_cfg3(dev, dtype, 1237, batch, per_img, ph)
    K = rois.shape[0]
    es = feats[0].element_size()
    shapes = [(batch, bench.C, h, w) for h, w in bench.SHAPES]
    out = torch.empty(K, bench.C, ph, ph, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last)
    _, lv = capi.msroi_align_forward(feats, rois, bench.SCALES, ph, ph, 2, False, out=out)
    gf = [torch.empty_like(f) for f in feats]
    fb = sum(f.numel() for f in feats) * es
    ob = out.numel() * es
    n = f"{tag}_{str(dtype).split('.')[-1]}_K{K}_p{ph}"
    ms, b = timed(lambda: capi.msroi_align_forward(feats, rois, bench.SCALES, ph, ph, 2, False, levels=lv, out=out))
    report(n + " fwd", ms, b, fb + ob)
    ms, b = timed(lambda: capi.msroi_align_backward(grad, rois, lv, bench.SCALES, shapes, ph, ph, 2, False,
                                                    layout=capi.NHWC, outs=gf))
    report(n + " bwd (memset incl.)", ms, b, fb + ob)
    ms, b = timed(lambda: capi.msroi_align_backward(grad, rois, lv, bench.SCALES, shapes, ph, ph, 2, False,
                                                    layout=capi.NHWC, outs=gf, zero_init=False))
    report(n + " bwd (no memset)", ms, b, fb + ob)


def cfg1():
    gen = torch.Generator(device=dev).manual_seed(1235)
    x = torch.randn(2, bench.C, 200, 272, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
    r = bench.make_rois(gen, 2, 1000, 16.0, 512.0, dev)
    g = torch.randn(2000, bench.C, 7, 7, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
    gi = torch.empty_like(x)
    nb = x.numel() * 4 + g.numel() * 4
    ms, b = timed(lambda: capi.roi_align_forward(x, r, 7, 7, 0.25, 2, True), iters=20)
    report("cfg1 fwd (L2 flushed)", ms, b, nb)
    ms, b = timed(lambda: capi.roi_align_backward(g, r, 0.25, 7, 7, 2, bench.C, 200, 272, 2, True, layout=capi.NHWC,
                                                  out=gi), iters=20)
    report("cfg1 bwd (L2 flushed)", ms, b, nb)


if __name__ == "__main__":
    legs = sys.argv[1:] or ["cfg3_fp32", "cfg3_bf16", "cfg1", "cfg4"]
    for leg in legs:
        if leg == "cfg3_fp32":
            cfg3(torch.float32)
        elif leg == "cfg3_bf16":
            cfg3(torch.bfloat16)
        elif leg == "cfg1":
            cfg1()
        elif leg == "cfg4":
            cfg3(torch.float32, batch=16, per_img=2000, ph=7, tag="cfg4slice")
