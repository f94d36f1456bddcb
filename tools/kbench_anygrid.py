#!/usr/bin/env python
"""Timing of the any-grid roi_align kernels against the generic kernels on the cfg3 / cfg1 shapes with adaptive
sampling (sampling_ratio 0) and sampling_ratio 3 (development aid).

    python tools/kbench_anygrid.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from cpu_vision_b200 import capi  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters=8):
    for _ in range(2):
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


def cfg3(dtype, sr, ph):
    feats, rois, grad = bench.make_cfg3(dev, dtype, 1237, 16, 512, ph)
    shapes = [(16, bench.C, h, w) for h, w in bench.SHAPES]
    lv = capi.fpn_level_map(rois, 2, 5)
    out = torch.empty(rois.shape[0], bench.C, ph, ph, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last)
    gf = [torch.empty_like(f) for f in feats]
    res = {}
    for sw in ("1", "0"):
        os.environ["B200DET_ANYGRID"] = sw
        f = timed(lambda: capi.msroi_align_forward(feats, rois, bench.SCALES, ph, ph, sr, False, levels=lv, out=out))
        b = timed(lambda: capi.msroi_align_backward(grad, rois, lv, bench.SCALES, shapes, ph, ph, sr, False,
                                                    layout=capi.NHWC, outs=gf))
        res[sw] = (f, b)
    print(f"cfg3 {str(dtype).split('.')[-1]} p{ph} sr={sr}: anygrid fwd {res['1'][0]:.3f} bwd {res['1'][1]:.3f} ms | "
          f"generic fwd {res['0'][0]:.3f} bwd {res['0'][1]:.3f} ms", flush=True)


def cfg1(sr):
    gen = torch.Generator(device=dev).manual_seed(1235)
    x = torch.randn(2, bench.C, 200, 272, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
    r = bench.make_rois(gen, 2, 1000, 16.0, 512.0, dev)
    g = torch.randn(2000, bench.C, 7, 7, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
    res = {}
    for sw in ("1", "0"):
        os.environ["B200DET_ANYGRID"] = sw
        f = timed(lambda: capi.roi_align_forward(x, r, 7, 7, 0.25, sr, True))
        b = timed(lambda: capi.roi_align_backward(g, r, 0.25, 7, 7, 2, bench.C, 200, 272, sr, True, layout=capi.NHWC))
        res[sw] = (f, b)
    print(f"cfg1 fp32 sr={sr}: anygrid fwd {res['1'][0]:.3f} bwd {res['1'][1]:.3f} ms | "
          f"generic fwd {res['0'][0]:.3f} bwd {res['0'][1]:.3f} ms", flush=True)


if __name__ == "__main__":
    for sr in (0, 3):
        cfg3(torch.float32, sr, 14)
    cfg3(torch.float32, 0, 7)
    cfg3(torch.bfloat16, 0, 14)
    cfg3(torch.float32, 2, 14)  # reference point: the sampling_ratio-2 kernels (the switch changes nothing here)
    for sr in (0, 2):
        cfg1(sr)
