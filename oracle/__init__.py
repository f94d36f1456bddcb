"""Parity oracle -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes bindings over
  * oracle/liboracle.so      the plain-C restatement (oracle/detection_oracle.c)
  * oracle/_ref/libtvref.so  the reference's own CPU kernels (oracle/ref_wrap.cpp)
operating on CPU torch tensors.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs import this package; nothing
under cpu_vision_b200/ does.

`impl="oracle"` selects the restatement, `impl="ref"` the compiled reference.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


def _build_if_needed():
    from . import build as _b

    _b.build_oracle()
    _b.build_ref()


def lib(impl="oracle"):
    if impl not in _libs:
        path = os.path.join(_HERE, "liboracle.so" if impl == "oracle" else "_ref/libtvref.so")
        if not os.path.exists(path):
            _build_if_needed()
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        _libs[impl] = C.CDLL(path)
        _declare(_libs[impl], impl)
    return _libs[impl]


def have_ref():
    try:
        lib("ref")
        return True
    except (OSError, FileNotFoundError):
        return False


_P = C.c_void_p
_I = C.c_int
_L = C.c_int64
_D = C.c_double


def _declare(l, impl):
    pre = "oracle" if impl == "oracle" else "tvref"
    long_t = C.c_long if impl == "oracle" else C.c_int64
    for sfx in ("f32", "f64"):
        f = getattr(l, f"{pre}_roi_align_forward_{sfx}")
        f.argtypes = [_P, _P, _P, _I, _I, _I, _I, _I, _I, _D, _I, _I]
        f.restype = None
        f = getattr(l, f"{pre}_roi_align_backward_{sfx}")
        f.argtypes = [_P, long_t, long_t, long_t, long_t, _P, _P, _I, _I, _I, _I, _I, _I, _D, _I, _I]
        f.restype = None
        f = getattr(l, f"{pre}_nms_{sfx}")
        f.argtypes = [_P, _P, _L, _D, _I, _P] if impl == "oracle" else [_P, _P, _L, _D, _P]
        f.restype = _L
        f = getattr(l, f"{pre}_roi_pool_forward_{sfx}")
        f.argtypes = [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _D]
        f.restype = None
        f = getattr(l, f"{pre}_roi_pool_backward_{sfx}")
        f.argtypes = [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, long_t, long_t, long_t, long_t]
        f.restype = None
        f = getattr(l, f"{pre}_ps_roi_pool_forward_{sfx}")
        f.argtypes = [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _D]
        f.restype = None
        f = getattr(l, f"{pre}_ps_roi_pool_backward_{sfx}")
        f.argtypes = [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _D]
        f.restype = None
        f = getattr(l, f"{pre}_ps_roi_align_forward_{sfx}")
        f.argtypes = [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _D, _I]
        f.restype = None
        f = getattr(l, f"{pre}_ps_roi_align_backward_{sfx}")
        f.argtypes = [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _D, _I]
        f.restype = None
    if impl == "oracle":
        for name in ("oracle_batched_nms_trick_f32", "oracle_batched_nms_vanilla_f32"):
            f = getattr(l, name)
            f.argtypes = [_P, _P, _P, _L, _D, _I, _P]
            f.restype = _L
        l.oracle_fpn_level_map_f32.argtypes = [_P, _L, _I, _I, _I, _I, _D, _P]
        l.oracle_fpn_level_map_f32.restype = None
        l.oracle_msroi_align_forward_f32.argtypes = [_P, _P, _P, _P, _I, _P, _P, _P, _I, _I, _I, _I, _I, _I]
        l.oracle_msroi_align_forward_f32.restype = None
        l.oracle_msroi_align_backward_f32.argtypes = [_P, C.c_long, C.c_long, C.c_long, C.c_long, _P, _P, _P,
                                                      _P, _P, _P, _I, _I, _I, _I, _I, _I, _I]
        l.oracle_msroi_align_backward_f32.restype = None
    else:
        l.tvref_roi_align_forward_f32_mt.argtypes = [_P, _P, _P, _I, _I, _I, _I, _I, _I, _D, _I, _I, _I]
        l.tvref_roi_align_forward_f32_mt.restype = None
        l.tvref_roi_align_backward_f32_mt.argtypes = [_P, _L, _L, _L, _L, _P, _P, _I, _I, _I, _I, _I, _I, _D,
                                                      _I, _I, _I]
        l.tvref_roi_align_backward_f32_mt.restype = _I
        l.tvref_set_num_threads.argtypes = [_I]


def _sfx(t):
    if t.dtype == torch.float32:
        return "f32"
    if t.dtype == torch.float64:
        return "f64"
    raise TypeError(f"oracle supports float32/float64, got {t.dtype}")


def _cpu(t):
    assert t.device.type == "cpu", "the oracle runs on CPU tensors"
    return t


def roi_align_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0, sampling_ratio=-1, aligned=False,
                      impl="oracle", threads=1):
    """x [N,C,H,W] (any strides; made NCHW-contiguous like the reference), rois [K,5]."""
    x = _cpu(x).contiguous()
    rois = _cpu(rois).to(x.dtype).contiguous()
    K, (N, Cc, H, W) = rois.shape[0], x.shape
    out = torch.zeros(K, Cc, pooled_h, pooled_w, dtype=x.dtype)
    if out.numel() == 0:
        return out
    l = lib(impl)
    pre = "oracle" if impl == "oracle" else "tvref"
    args = [x.data_ptr(), rois.data_ptr(), out.data_ptr(), K, Cc, H, W, pooled_h, pooled_w,
            float(spatial_scale), int(sampling_ratio), int(bool(aligned))]
    if impl == "ref" and threads > 1 and x.dtype == torch.float32:
        l.tvref_roi_align_forward_f32_mt(*args, int(threads))
    else:
        getattr(l, f"{pre}_roi_align_forward_{_sfx(x)}")(*args)
    return out


def roi_align_backward(grad, rois, spatial_scale, pooled_h, pooled_w, N, Cc, H, W, sampling_ratio, aligned,
                       impl="oracle", threads=1):
    """grad [K,C,PH,PW] with arbitrary strides -> grad_input [N,C,H,W] NCHW."""
    grad = _cpu(grad)
    rois = _cpu(rois).to(grad.dtype).contiguous()
    gi = torch.zeros(N, Cc, H, W, dtype=grad.dtype)
    if grad.numel() == 0:
        return gi
    l = lib(impl)
    pre = "oracle" if impl == "oracle" else "tvref"
    s = grad.stride()
    args = [grad.data_ptr(), s[0], s[1], s[2], s[3], rois.data_ptr(), gi.data_ptr(), rois.shape[0], Cc, H, W,
            pooled_h, pooled_w, float(spatial_scale), int(sampling_ratio), int(bool(aligned))]
    if impl == "ref" and threads > 1 and grad.dtype == torch.float32:
        used = l.tvref_roi_align_backward_f32_mt(*args, int(threads))
        if used == 0:
            getattr(l, f"{pre}_roi_align_backward_{_sfx(grad)}")(*args)
    else:
        getattr(l, f"{pre}_roi_align_backward_{_sfx(grad)}")(*args)
    return gi


def nms(boxes, scores, iou_threshold, thr_is_float=True, impl="oracle"):
    boxes = _cpu(boxes).contiguous()
    scores = _cpu(scores).to(boxes.dtype).contiguous()
    n = boxes.shape[0]
    keep = torch.empty(max(n, 1), dtype=torch.int64)
    l = lib(impl)
    if impl == "oracle":
        k = getattr(l, f"oracle_nms_{_sfx(boxes)}")(boxes.data_ptr(), scores.data_ptr(), n, float(iou_threshold),
                                                     int(thr_is_float), keep.data_ptr())
    else:
        k = getattr(l, f"tvref_nms_{_sfx(boxes)}")(boxes.data_ptr(), scores.data_ptr(), n, float(iou_threshold),
                                                    keep.data_ptr())
    return keep[:k].clone()


def batched_nms(boxes, scores, idxs, iou_threshold, mode="auto_cuda", thr_is_float=True):
    """mode: 'trick' (boxes.py:78-95), 'vanilla' (boxes.py:98-112) or 'auto_cuda'
    = the branch the reference takes for a CUDA tensor (numel > 20000 -> vanilla)."""
    boxes = _cpu(boxes).float().contiguous()
    scores = _cpu(scores).float().contiguous()
    idxs = _cpu(idxs).to(torch.int64).contiguous()
    n = boxes.shape[0]
    if mode == "auto_cuda":
        mode = "vanilla" if boxes.numel() > 20000 else "trick"
    keep = torch.empty(max(n, 1), dtype=torch.int64)
    f = lib().oracle_batched_nms_trick_f32 if mode == "trick" else lib().oracle_batched_nms_vanilla_f32
    k = f(boxes.data_ptr(), scores.data_ptr(), idxs.data_ptr(), n, float(iou_threshold), int(thr_is_float),
          keep.data_ptr())
    return keep[:k].clone()


def fpn_level_map(rois, k_min, k_max, canonical_scale=224, canonical_level=4, eps=1e-6):
    rois = _cpu(rois).float().contiguous()
    out = torch.empty(rois.shape[0], dtype=torch.int64)
    lib().oracle_fpn_level_map_f32(rois.data_ptr(), rois.shape[0], k_min, k_max, canonical_scale,
                                   canonical_level, float(eps), out.data_ptr())
    return out


def _level_arrays(feats, scales):
    L = len(feats)
    ptrs = (C.c_void_p * L)(*[f.data_ptr() for f in feats])
    hs = (C.c_int * L)(*[f.shape[2] for f in feats])
    ws = (C.c_int * L)(*[f.shape[3] for f in feats])
    sc = (C.c_double * L)(*[float(s) for s in scales])
    return ptrs, hs, ws, sc


def msroi_align_forward(feats, rois, levels, scales, pooled_h, pooled_w, sampling_ratio, aligned=False):
    feats = [_cpu(f).float().contiguous() for f in feats]
    rois = _cpu(rois).float().contiguous()
    levels = _cpu(levels).to(torch.int64).contiguous()
    K, Cc = rois.shape[0], feats[0].shape[1]
    out = torch.zeros(K, Cc, pooled_h, pooled_w)
    ptrs, hs, ws, sc = _level_arrays(feats, scales)
    lib().oracle_msroi_align_forward_f32(ptrs, hs, ws, sc, len(feats), rois.data_ptr(), levels.data_ptr(),
                                         out.data_ptr(), K, Cc, pooled_h, pooled_w, int(sampling_ratio),
                                         int(bool(aligned)))
    return out


def msroi_align_backward(grad, rois, levels, scales, shapes, pooled_h, pooled_w, sampling_ratio, aligned=False):
    """shapes: list of (N, C, H, W) per level -> list of fp32 NCHW gradients."""
    grad = _cpu(grad).float()
    rois = _cpu(rois).float().contiguous()
    levels = _cpu(levels).to(torch.int64).contiguous()
    outs = [torch.zeros(*s) for s in shapes]
    ptrs, hs, ws, sc = _level_arrays(outs, scales)
    s = grad.stride()
    lib().oracle_msroi_align_backward_f32(grad.data_ptr(), s[0], s[1], s[2], s[3], rois.data_ptr(),
                                          levels.data_ptr(), ptrs, hs, ws, sc, len(outs), rois.shape[0],
                                          shapes[0][1], pooled_h, pooled_w, int(sampling_ratio),
                                          int(bool(aligned)))
    return outs


def roi_pool_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0, impl="oracle"):
    """cpu/roi_pool_kernel.cpp:127-186 -> (output, argmax int32), NCHW."""
    x = _cpu(x).contiguous()
    rois = _cpu(rois).to(x.dtype).contiguous()
    N, Cc, H, W = x.shape
    K = rois.shape[0]
    out = torch.zeros(K, Cc, pooled_h, pooled_w, dtype=x.dtype)
    am = torch.zeros(K, Cc, pooled_h, pooled_w, dtype=torch.int32)
    pre = "oracle" if impl == "oracle" else "tvref"
    if out.numel():
        getattr(lib(impl), f"{pre}_roi_pool_forward_{_sfx(x)}")(x.data_ptr(), rois.data_ptr(), out.data_ptr(),
                                                                am.data_ptr(), K, Cc, H, W, pooled_h, pooled_w,
                                                                float(spatial_scale))
    return out, am


def roi_pool_backward(grad, rois, argmax, pooled_h, pooled_w, N, Cc, H, W, impl="oracle"):
    """cpu/roi_pool_kernel.cpp:188-249 (grad strides honoured)."""
    grad = _cpu(grad)
    rois = _cpu(rois).to(grad.dtype).contiguous()
    am = _cpu(argmax).to(torch.int32).contiguous()
    gi = torch.zeros(N, Cc, H, W, dtype=grad.dtype)
    pre = "oracle" if impl == "oracle" else "tvref"
    if grad.numel():
        getattr(lib(impl), f"{pre}_roi_pool_backward_{_sfx(grad)}")(grad.data_ptr(), am.data_ptr(), rois.data_ptr(),
                                                                    gi.data_ptr(), rois.shape[0], Cc, H, W, pooled_h,
                                                                    pooled_w, *grad.stride())
    return gi


def ps_roi_pool_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0, impl="oracle"):
    """cpu/ps_roi_pool_kernel.cpp:148-205 -> (output [K, C/(PH*PW), PH, PW], channel_mapping int32), NCHW."""
    x = _cpu(x).contiguous()
    rois = _cpu(rois).to(x.dtype).contiguous()
    N, Cc, H, W = x.shape
    K = rois.shape[0]
    co = Cc // (pooled_h * pooled_w)
    out = torch.zeros(K, co, pooled_h, pooled_w, dtype=x.dtype)
    cm = torch.zeros(K, co, pooled_h, pooled_w, dtype=torch.int32)
    pre = "oracle" if impl == "oracle" else "tvref"
    if out.numel():
        getattr(lib(impl), f"{pre}_ps_roi_pool_forward_{_sfx(x)}")(x.data_ptr(), rois.data_ptr(), out.data_ptr(),
                                                                   cm.data_ptr(), K, Cc, H, W, pooled_h, pooled_w,
                                                                   float(spatial_scale))
    return out, cm


def ps_roi_pool_backward(grad, rois, mapping, spatial_scale, pooled_h, pooled_w, N, Cc, H, W, impl="oracle"):
    """cpu/ps_roi_pool_kernel.cpp:207-260 (grad.contiguous() like the reference)."""
    grad = _cpu(grad).contiguous()
    rois = _cpu(rois).to(grad.dtype).contiguous()
    cm = _cpu(mapping).to(torch.int32).contiguous()
    gi = torch.zeros(N, Cc, H, W, dtype=grad.dtype)
    pre = "oracle" if impl == "oracle" else "tvref"
    if grad.numel():
        getattr(lib(impl), f"{pre}_ps_roi_pool_backward_{_sfx(grad)}")(grad.data_ptr(), cm.data_ptr(), rois.data_ptr(),
                                                                       gi.data_ptr(), rois.shape[0], Cc, H, W,
                                                                       pooled_h, pooled_w, float(spatial_scale))
    return gi


def ps_roi_align_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0, sampling_ratio=-1, impl="oracle"):
    """cpu/ps_roi_align_kernel.cpp:302-365 -> (output [K, C/(PH*PW), PH, PW], channel_mapping int32), NCHW."""
    x = _cpu(x).contiguous()
    rois = _cpu(rois).to(x.dtype).contiguous()
    N, Cc, H, W = x.shape
    K = rois.shape[0]
    co = Cc // (pooled_h * pooled_w)
    out = torch.zeros(K, co, pooled_h, pooled_w, dtype=x.dtype)
    cm = torch.zeros(K, co, pooled_h, pooled_w, dtype=torch.int32)
    pre = "oracle" if impl == "oracle" else "tvref"
    if out.numel():
        getattr(lib(impl), f"{pre}_ps_roi_align_forward_{_sfx(x)}")(x.data_ptr(), rois.data_ptr(), out.data_ptr(),
                                                                    cm.data_ptr(), K, Cc, H, W, pooled_h, pooled_w,
                                                                    float(spatial_scale), int(sampling_ratio))
    return out, cm


def ps_roi_align_backward(grad, rois, mapping, spatial_scale, pooled_h, pooled_w, sampling_ratio, N, Cc, H, W,
                          impl="oracle"):
    """cpu/ps_roi_align_kernel.cpp:367-431 (grad.contiguous() like the reference)."""
    grad = _cpu(grad).contiguous()
    rois = _cpu(rois).to(grad.dtype).contiguous()
    cm = _cpu(mapping).to(torch.int32).contiguous()
    gi = torch.zeros(N, Cc, H, W, dtype=grad.dtype)
    pre = "oracle" if impl == "oracle" else "tvref"
    if grad.numel():
        getattr(lib(impl), f"{pre}_ps_roi_align_backward_{_sfx(grad)}")(grad.data_ptr(), cm.data_ptr(), rois.data_ptr(),
                                                                        gi.data_ptr(), rois.shape[0], Cc, H, W,
                                                                        pooled_h, pooled_w, float(spatial_scale),
                                                                        int(sampling_ratio))
    return gi
