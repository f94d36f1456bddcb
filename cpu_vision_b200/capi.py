"""ctypes binding of the C-ABI (include/b200_detection_ops.h) on torch tensors.

This is the same boundary a non-Python host would bind; the
parity tests and bench.py go through it so that what is measured is the C-ABI
library and not the PyTorch dispatcher.  Device entry points take CUDA tensors
and launch on torch's current stream; `*_host` entry points take CPU tensors
(pinned for full PCIe rate) and return after the results are back on the host.
"""
import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib

F32, F16, BF16, F64 = 0, 1, 2, 3
NCHW, NHWC = 0, 1
BNMS_AUTO, BNMS_TRICK, BNMS_VANILLA = -1, 0, 1

_DT = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16, torch.float64: F64}
_P, _I, _L, _D, _S = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_size_t

SYMBOLS = [
    "b200det_version", "b200det_last_error", "b200det_roi_align_forward", "b200det_roi_align_backward",
    "b200det_fpn_level_map", "b200det_msroi_align_forward", "b200det_msroi_align_backward",
    "b200det_nms_workspace_bytes", "b200det_nms", "b200det_batched_nms", "b200det_roi_align_forward_host",
    "b200det_roi_align_backward_host", "b200det_msroi_align_forward_host", "b200det_msroi_align_backward_host",
    "b200det_nms_host", "b200det_batched_nms_host", "b200det_host_release",
    "b200det_nms_segmented_workspace_bytes", "b200det_nms_segmented", "b200det_clip_filter_boxes", "b200det_relayout",
    "b200det_roi_pool_forward", "b200det_roi_pool_backward", "b200det_ps_roi_pool_forward",
    "b200det_ps_roi_pool_backward", "b200det_ps_roi_align_forward", "b200det_ps_roi_align_backward",
    "b200det_rpn_select", "b200det_roi_heads_decode", "b200det_nms_static",
    "b200det_msroi_align_forward_push", "b200det_peer_alloc", "b200det_peer_free", "b200det_peer_export",
    "b200det_peer_open", "b200det_peer_close", "b200det_segmented_topk",
]

_declared = False


def lib():
    global _declared
    l = _lib.capi()
    if not _declared:
        l.b200det_version.restype = _I
        l.b200det_last_error.restype = C.c_char_p
        l.b200det_roi_align_forward.argtypes = [_P, _I, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _I, _I, _P]
        l.b200det_roi_align_backward.argtypes = [_P, _P, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _I, _I, _I, _P]
        l.b200det_fpn_level_map.argtypes = [_P, _I, _L, _I, _I, _I, _I, _D, _P, _P]
        l.b200det_msroi_align_forward.argtypes = [_P, _P, _P, _P, _I, _I, _P, _P, _I, _P, _I, _I, _L, _L, _L, _L, _L,
                                                  _I, _I, _I, _I, _I, _I, _D, _P]
        l.b200det_msroi_align_backward.argtypes = [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _L, _L, _L, _L, _L,
                                                   _I, _I, _I, _P]
        l.b200det_msroi_align_forward_push.argtypes = [_P, _P, _P, _P, _I, _I, _P, _P, _I, _P, _P, _I, _I, _I, _L, _L,
                                                       _L, _L, _L, _I, _I, _I, _I, _I, _I, _D, _P]
        l.b200det_segmented_topk.argtypes = [_P, _L, _L, _P, _I, _L, _P, _L, _P]
        l.b200det_peer_alloc.argtypes = [_S, _P]
        l.b200det_peer_free.argtypes = [_P]
        l.b200det_peer_export.argtypes = [_P, _P]
        l.b200det_peer_open.argtypes = [_P, _P]
        l.b200det_peer_close.argtypes = [_P]
        l.b200det_nms_workspace_bytes.argtypes = [_L]
        l.b200det_nms_workspace_bytes.restype = _S
        l.b200det_nms.argtypes = [_P, _P, _I, _L, _D, _P, _P, _P, _S, _P]
        l.b200det_batched_nms.argtypes = [_P, _P, _P, _I, _L, _D, _I, _P, _P, _P, _S, _P]
        l.b200det_nms_static.argtypes = [_P, _P, _I, _L, _D, _P, _P, _P]
        l.b200det_roi_align_forward_host.argtypes = [_P, _I, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _I, _I, _I]
        l.b200det_roi_align_backward_host.argtypes = [_P, _P, _L, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _I,
                                                      _I, _I]
        l.b200det_msroi_align_forward_host.argtypes = [_P, _P, _P, _P, _I, _I, _P, _P, _P, _I, _I, _L, _L, _L, _L, _L,
                                                       _I, _I, _I, _I, _I, _I, _D, _I]
        l.b200det_msroi_align_backward_host.argtypes = [_P, _P, _L, _P, _P, _P, _P, _P, _P, _I, _I, _I, _L, _L, _L,
                                                        _L, _L, _I, _I, _I]
        l.b200det_nms_host.argtypes = [_P, _P, _I, _L, _D, _P, _P, _I]
        l.b200det_batched_nms_host.argtypes = [_P, _P, _P, _I, _L, _D, _I, _P, _P, _I]
        l.b200det_host_release.restype = None
        l.b200det_nms_segmented_workspace_bytes.argtypes = [_L, _P, _L]
        l.b200det_nms_segmented_workspace_bytes.restype = _S
        l.b200det_nms_segmented.argtypes = [_P, _P, _P, _P, _I, _L, _P, _L, _D, _I, _L, _P, _P, _P, _S, _P]
        l.b200det_clip_filter_boxes.argtypes = [_P, _P, _I, _L, _P, _L, _P, _D, _D, _I, _P, _P, _P]
        l.b200det_relayout.argtypes = [_P, _P, _I, _L, _L, _L, _P]
        l.b200det_rpn_select.argtypes = [_P, _P, _P, _L, _L, _L, _P, _P, _I, _D, _D, _P, _P, _P, _P, _P]
        l.b200det_roi_heads_decode.argtypes = [_P, _P, _P, _P, _P, _L, _L, _P, _D, _D, _D, _P, _P, _P, _P]
        l.b200det_roi_pool_forward.argtypes = [_P, _I, _P, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _P]
        l.b200det_roi_pool_backward.argtypes = [_P, _P, _P, _I, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _P]
        l.b200det_ps_roi_pool_forward.argtypes = [_P, _I, _P, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _P]
        l.b200det_ps_roi_pool_backward.argtypes = [_P, _P, _P, _I, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _P]
        l.b200det_ps_roi_align_forward.argtypes = [_P, _I, _P, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _I, _P]
        l.b200det_ps_roi_align_backward.argtypes = [_P, _P, _P, _I, _P, _P, _I, _I, _L, _L, _L, _L, _L, _L, _L, _D, _I,
                                                    _P]
        _declared = True
    return l


class B200DetError(RuntimeError):
    pass


def _check(rc, who):
    if rc != 0:
        raise B200DetError(f"{who} failed ({rc}): {lib().b200det_last_error().decode()}")


def _dt(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}") from None


def _layout_of(t):
    """(layout, dense tensor) for a logically [N,C,H,W] tensor."""
    if t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last) and not t.is_contiguous():
        return NHWC, t
    return NCHW, t.contiguous()


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _arith(rois, feat_dtype):
    return rois.to(torch.float64 if feat_dtype == torch.float64 else torch.float32).contiguous()


def _empty_like_layout(shape, like, layout):
    mf = torch.channels_last if layout == NHWC else torch.contiguous_format
    return torch.empty(shape, dtype=like.dtype, device=like.device, memory_format=mf)


def _strides(t):
    return (_L * 4)(*t.stride())


# --------------------------------------------------------------------------- device entry points
def roi_align_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0, sampling_ratio=-1, aligned=False,
                      out_layout=None):
    assert x.is_cuda and rois.is_cuda
    layout, x = _layout_of(x)
    out_layout = layout if out_layout is None else out_layout
    r = _arith(rois, x.dtype)
    N, Cc, H, W = x.shape
    K = r.shape[0]
    out = _empty_like_layout((K, Cc, pooled_h, pooled_w), x, out_layout)
    with torch.cuda.device(x.device):
        _check(lib().b200det_roi_align_forward(x.data_ptr(), layout, r.data_ptr(), out.data_ptr(), out_layout, _dt(x),
                                               N, Cc, H, W, K, pooled_h, pooled_w, float(spatial_scale),
                                               int(sampling_ratio), int(bool(aligned)), _stream(x)),
               "b200det_roi_align_forward")
    return out


def roi_align_backward(grad, rois, spatial_scale, pooled_h, pooled_w, N, Cc, H, W, sampling_ratio, aligned,
                       layout=NCHW, out=None, zero_init=True, deterministic=False):
    assert grad.is_cuda and rois.is_cuda
    r = _arith(rois, grad.dtype)
    gi = out if out is not None else _empty_like_layout((N, Cc, H, W), grad, layout)
    with torch.cuda.device(grad.device):
        _check(lib().b200det_roi_align_backward(grad.data_ptr(), _strides(grad), r.data_ptr(), gi.data_ptr(), layout,
                                                _dt(grad), N, Cc, H, W, r.shape[0], pooled_h, pooled_w,
                                                float(spatial_scale), int(sampling_ratio), int(bool(aligned)),
                                                int(bool(zero_init)) | (2 if deterministic else 0), _stream(grad)),
               "b200det_roi_align_backward")
    return gi


def fpn_level_map(rois, k_min, k_max, canonical_scale=224, canonical_level=4, eps=1e-6):
    r = rois.to(torch.float32).contiguous()
    lv = torch.empty(r.shape[0], dtype=torch.int32, device=r.device)
    with torch.cuda.device(r.device):
        _check(lib().b200det_fpn_level_map(r.data_ptr(), F32, r.shape[0], k_min, k_max, canonical_scale,
                                           canonical_level, float(eps), lv.data_ptr(), _stream(r)),
               "b200det_fpn_level_map")
    return lv


def _level_arrays(feats: Sequence[torch.Tensor], scales):
    L = len(feats)
    return ((_P * L)(*[f.data_ptr() for f in feats]), (_L * L)(*[f.shape[2] for f in feats]),
            (_L * L)(*[f.shape[3] for f in feats]), (_D * L)(*[float(s) for s in scales]))


def msroi_align_forward(feats: List[torch.Tensor], rois, scales, pooled_h, pooled_w, sampling_ratio, aligned=False,
                        k_min=2, k_max=5, canonical_scale=224, canonical_level=4, eps=1e-6,
                        levels: Optional[torch.Tensor] = None, out=None):
    """Returns (output [K,C,PH,PW], levels int32 [K])."""
    layout, f0 = _layout_of(feats[0])
    mf = torch.channels_last if layout == NHWC else torch.contiguous_format
    feats = [f.contiguous(memory_format=mf) for f in feats]
    r = _arith(rois, f0.dtype)
    N, Cc = f0.shape[:2]
    K = r.shape[0]
    if out is None:
        out = _empty_like_layout((K, Cc, pooled_h, pooled_w), f0, layout)
    given = levels is not None
    if not given:
        levels = torch.empty(K, dtype=torch.int32, device=f0.device)
    ptrs, hs, ws, sc = _level_arrays(feats, scales)
    with torch.cuda.device(f0.device):
        _check(lib().b200det_msroi_align_forward(ptrs, hs, ws, sc, len(feats), layout, r.data_ptr(), levels.data_ptr(),
                                                 int(given), out.data_ptr(), layout, _dt(f0), N, Cc, K, pooled_h,
                                                 pooled_w, int(sampling_ratio), int(bool(aligned)), k_min, k_max,
                                                 canonical_scale, canonical_level, float(eps), _stream(f0)),
               "b200det_msroi_align_forward")
    return out, levels


def msroi_align_forward_push(feats: List[torch.Tensor], rois, scales, pooled_h, pooled_w, sampling_ratio, out,
                             peer_ptrs, aligned=False, k_min=2, k_max=5, canonical_scale=224, canonical_level=4,
                             eps=1e-6, levels: Optional[torch.Tensor] = None):
    """MultiScaleRoIAlign forward fused with the push all-gather: every pooled row is stored to `out` (this rank's
    slice of its own gather buffer) and to the same row behind every pointer of `peer_ptrs` (this rank's slice of
    the peers' buffers, device addresses as ints; sharding.PushGather owns them).  Returns (out, levels)."""
    layout, f0 = _layout_of(feats[0])
    if layout != NHWC:
        raise ValueError("msroi_align_forward_push: channels_last features only")
    feats = [f.contiguous(memory_format=torch.channels_last) for f in feats]
    r = _arith(rois, f0.dtype)
    N, Cc = f0.shape[:2]
    K = r.shape[0]
    if tuple(out.shape) != (K, Cc, pooled_h, pooled_w) or not out.is_contiguous(memory_format=torch.channels_last):
        raise ValueError("msroi_align_forward_push: `out` must be a channels_last [K, C, PH, PW] tensor")
    given = levels is not None
    if not given:
        levels = torch.empty(K, dtype=torch.int32, device=f0.device)
    ptrs, hs, ws, sc = _level_arrays(feats, scales)
    peers = (C.c_void_p * max(len(peer_ptrs), 1))(*[int(p) for p in peer_ptrs])
    with torch.cuda.device(f0.device):
        _check(lib().b200det_msroi_align_forward_push(ptrs, hs, ws, sc, len(feats), layout, r.data_ptr(),
                                                      levels.data_ptr(), int(given), out.data_ptr(), peers,
                                                      len(peer_ptrs), layout, _dt(f0), N, Cc, K, pooled_h, pooled_w,
                                                      int(sampling_ratio), int(bool(aligned)), k_min, k_max,
                                                      canonical_scale, canonical_level, float(eps), _stream(f0)),
               "b200det_msroi_align_forward_push")
    return out, levels


SEGMENTED_TOPK_MAX_K = 2048


def segmented_topk(scores, num_anchors_per_level, k):
    """scores fp32 [B, A] -> int64 [B, sum(min(k, n_l))]: per level the indices (into A) of its top-k scores in
    descending order, ties by ascending index (b200det_segmented_topk; rpn.py:231-240 in one launch)."""
    s = scores.contiguous()
    B, A = s.shape
    ends, acc = [], 0
    for n in num_anchors_per_level:
        acc += int(n)
        ends.append(acc)
    P = sum(min(int(k), int(n)) for n in num_anchors_per_level)
    out = torch.empty((B, P), dtype=torch.int64, device=s.device)
    le = (C.c_int32 * len(ends))(*ends)
    with torch.cuda.device(s.device):
        _check(lib().b200det_segmented_topk(s.data_ptr(), B, A, le, len(ends), int(k), out.data_ptr(), P, _stream(s)),
               "b200det_segmented_topk")
    return out


def peer_alloc(nbytes: int, device) -> int:
    """A cudaMalloc of its own on `device` (so that an IPC handle covers exactly it); returns the address."""
    p = C.c_void_p(0)
    with torch.cuda.device(device):
        _check(lib().b200det_peer_alloc(int(nbytes), C.byref(p)), "b200det_peer_alloc")
    return int(p.value)


def peer_free(ptr: int, device):
    with torch.cuda.device(device):
        _check(lib().b200det_peer_free(C.c_void_p(ptr)), "b200det_peer_free")


def peer_export(ptr: int, device) -> bytes:
    h = C.create_string_buffer(64)
    with torch.cuda.device(device):
        _check(lib().b200det_peer_export(C.c_void_p(ptr), h), "b200det_peer_export")
    return h.raw


def peer_open(handle: bytes, device) -> int:
    p = C.c_void_p(0)
    h = C.create_string_buffer(handle, 64)
    with torch.cuda.device(device):
        _check(lib().b200det_peer_open(h, C.byref(p)), "b200det_peer_open")
    return int(p.value)


def peer_close(ptr: int, device):
    with torch.cuda.device(device):
        _check(lib().b200det_peer_close(C.c_void_p(ptr)), "b200det_peer_close")


def msroi_align_backward(grad, rois, levels, scales, shapes, pooled_h, pooled_w, sampling_ratio, aligned=False,
                         layout=NCHW, outs=None, zero_init=True, deterministic=False):
    """shapes: [(N,C,H,W)] per level -> list of gradients in `layout`."""
    r = _arith(rois, grad.dtype)
    if outs is None:
        outs = [_empty_like_layout(tuple(s), grad, layout) for s in shapes]
    ptrs, hs, ws, sc = _level_arrays(outs, scales)
    N, Cc = shapes[0][:2]
    with torch.cuda.device(grad.device):
        _check(lib().b200det_msroi_align_backward(grad.data_ptr(), _strides(grad), r.data_ptr(), levels.data_ptr(),
                                                  ptrs, hs, ws, sc, len(outs), layout, _dt(grad), N, Cc, r.shape[0],
                                                  pooled_h, pooled_w, int(sampling_ratio), int(bool(aligned)),
                                                  int(bool(zero_init)) | (2 if deterministic else 0), _stream(grad)),
               "b200det_msroi_align_backward")
    return outs


class NmsWorkspace:
    """Reusable device scratch + output buffers for N boxes (no allocation per call)."""

    def __init__(self, n, device):
        self.n = n
        self.bytes = lib().b200det_nms_workspace_bytes(n)
        self.ws = torch.empty(self.bytes, dtype=torch.uint8, device=device)
        self.keep = torch.empty(n + 1, dtype=torch.int64, device=device)


def nms(boxes, scores, iou_threshold, workspace: Optional[NmsWorkspace] = None, sync=True):
    """Returns kept indices (sync=True) or (keep_buffer, count_tensor) still on the device."""
    boxes = boxes.contiguous()
    scores = scores.to(boxes.dtype).contiguous()
    n = boxes.shape[0]
    w = workspace if workspace is not None and workspace.n >= n else NmsWorkspace(n, boxes.device)
    cnt = w.keep[w.n:]
    with torch.cuda.device(boxes.device):
        _check(lib().b200det_nms(boxes.data_ptr(), scores.data_ptr(), _dt(boxes), n, float(iou_threshold),
                                 w.keep.data_ptr(), cnt.data_ptr(), w.ws.data_ptr(), w.bytes, _stream(boxes)),
               "b200det_nms")
    if not sync:
        return w.keep, cnt
    return w.keep[: int(cnt.item())].clone()


NMS_STATIC_MAX_BOXES = 8192  # B200DET_NMS_STATIC_MAX_BOXES (include/b200_detection_ops.h)


def nms_static(boxes, scores, iou_threshold, keep, num_keep):
    """b200det_nms_static: NMS on the library's built-in scratch array -- no allocation inside the call and no
    workspace; the caller owns the outputs (`keep` int64 [>= N], `num_keep` int64 [1], both on the device) and
    the inputs must already be contiguous and of one dtype.  At most NMS_STATIC_MAX_BOXES boxes; calls that
    overlap on the GPU race on the shared scratch.  Asynchronous: nothing is read back here."""
    n = boxes.shape[0]
    if n and not (boxes.is_contiguous() and scores.is_contiguous() and scores.dtype == boxes.dtype):
        raise ValueError("nms_static: contiguous boxes / scores of one dtype required (the call does not copy)")
    if keep.dtype != torch.int64 or num_keep.dtype != torch.int64 or keep.numel() < n or num_keep.numel() < 1:
        raise ValueError("nms_static: keep must be int64 [>= N] and num_keep int64 [1]")
    with torch.cuda.device(boxes.device):
        _check(lib().b200det_nms_static(boxes.data_ptr(), scores.data_ptr(), _dt(boxes), n, float(iou_threshold),
                                        keep.data_ptr(), num_keep.data_ptr(), _stream(boxes)),
               "b200det_nms_static")


def batched_nms(boxes, scores, idxs, iou_threshold, mode=BNMS_AUTO, workspace: Optional[NmsWorkspace] = None,
                sync=True):
    boxes = boxes.contiguous()
    scores = scores.to(boxes.dtype).contiguous()
    idxs = idxs.to(torch.int64).contiguous()
    n = boxes.shape[0]
    w = workspace if workspace is not None and workspace.n >= n else NmsWorkspace(n, boxes.device)
    cnt = w.keep[w.n:]
    with torch.cuda.device(boxes.device):
        _check(lib().b200det_batched_nms(boxes.data_ptr(), scores.data_ptr(), idxs.data_ptr(), _dt(boxes), n,
                                         float(iou_threshold), int(mode), w.keep.data_ptr(), cnt.data_ptr(),
                                         w.ws.data_ptr(), w.bytes, _stream(boxes)),
               "b200det_batched_nms")
    if not sync:
        return w.keep, cnt
    return w.keep[: int(cnt.item())].clone()


def nms_segmented(boxes, scores, idxs, seg_offsets: Sequence[int], iou_threshold, valid=None, mode=BNMS_AUTO,
                  top_n=0, sync=True):
    """Per-image class-aware NMS over S box sets in one launch sequence (b200det_nms_segmented).

    seg_offsets: S + 1 host integers.  Returns a list of S index tensors (sync=True; ONE read-back of
    the S counts) or (keep, counts) still on the device: set s owns keep[off[s] : off[s] + counts[s]].
    Indices are positions in the concatenated `boxes`."""
    boxes = boxes.contiguous()
    scores = scores.to(boxes.dtype).contiguous()
    n = boxes.shape[0]
    off = torch.as_tensor(list(seg_offsets), dtype=torch.int64)
    S = off.numel() - 1
    if idxs is not None:
        idxs = idxs.to(torch.int64).contiguous()
    if valid is not None:
        valid = valid.to(torch.uint8).contiguous()
    keep = torch.empty(max(n, 1), dtype=torch.int64, device=boxes.device)
    counts = torch.zeros(max(S, 1), dtype=torch.int64, device=boxes.device)
    if S > 0:
        nbytes = lib().b200det_nms_segmented_workspace_bytes(n, off.data_ptr(), S)
        if nbytes == 0:
            raise B200DetError("nms_segmented: seg_offsets must be a non-decreasing partition of [0, N)")
        ws = torch.empty(nbytes, dtype=torch.uint8, device=boxes.device)
        with torch.cuda.device(boxes.device):
            _check(lib().b200det_nms_segmented(boxes.data_ptr(), scores.data_ptr(),
                                               idxs.data_ptr() if idxs is not None else None,
                                               valid.data_ptr() if valid is not None else None, _dt(boxes), n,
                                               off.data_ptr(), S, float(iou_threshold), int(mode), int(top_n),
                                               keep.data_ptr(), counts.data_ptr(), ws.data_ptr(), nbytes,
                                               _stream(boxes)), "b200det_nms_segmented")
    if not sync:
        return keep, counts[:S]
    c = counts[:S].tolist()
    o = off.tolist()
    return [keep[o[s]: o[s] + c[s]] for s in range(S)]


def clip_filter_boxes(boxes, scores, seg_offsets: Sequence[int], image_shapes, min_size, score_thresh, strict):
    """Clipped boxes + validity byte per box (b200det_clip_filter_boxes); image_shapes: S (height, width)."""
    boxes = boxes.contiguous()
    scores = scores.to(boxes.dtype).contiguous()
    n = boxes.shape[0]
    off = torch.tensor(list(seg_offsets), dtype=torch.int32, device=boxes.device)
    sizes = torch.tensor([[float(h), float(w)] for h, w in image_shapes], dtype=torch.float32, device=boxes.device)
    out = torch.empty_like(boxes)
    valid = torch.empty(max(n, 1), dtype=torch.uint8, device=boxes.device)
    with torch.cuda.device(boxes.device):
        _check(lib().b200det_clip_filter_boxes(boxes.data_ptr(), scores.data_ptr(), _dt(boxes), n, off.data_ptr(),
                                               off.numel() - 1, sizes.data_ptr(), float(min_size), float(score_thresh),
                                               int(bool(strict)), out.data_ptr(), valid.data_ptr(), _stream(boxes)),
               "b200det_clip_filter_boxes")
    return out, valid[:n]


def rpn_select(proposals, objectness, top_idx, image_shapes, num_anchors_per_level, min_size, score_thresh):
    """b200det_rpn_select: [B, A, 4] fp32 proposals, [B, A] logits, int64 [B, P] top-k indices ->
    (boxes [B*P, 4], scores [B*P], levels int64 [B*P], valid uint8 [B*P])."""
    proposals, objectness, top_idx = proposals.contiguous(), objectness.contiguous(), top_idx.contiguous()
    B, A = objectness.shape
    P = top_idx.shape[1]
    dev = proposals.device
    sizes = torch.tensor([[float(h), float(w)] for h, w in image_shapes], dtype=torch.float32, device=dev)
    ends, acc = [], 0
    for n in num_anchors_per_level:
        acc += int(n)
        ends.append(acc)
    le = (C.c_int32 * len(ends))(*ends)
    boxes = torch.empty(B * P, 4, dtype=torch.float32, device=dev)
    scores = torch.empty(B * P, dtype=torch.float32, device=dev)
    levels = torch.empty(B * P, dtype=torch.int64, device=dev)
    valid = torch.empty(max(B * P, 1), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _check(lib().b200det_rpn_select(proposals.data_ptr(), objectness.data_ptr(), top_idx.data_ptr(), B, A, P,
                                        sizes.data_ptr(), le, len(ends), float(min_size), float(score_thresh),
                                        boxes.data_ptr(), scores.data_ptr(), levels.data_ptr(), valid.data_ptr(),
                                        _stream(proposals)), "b200det_rpn_select")
    return boxes, scores, levels, valid[:B * P]


def roi_heads_decode(class_logits, box_regression, proposals, image_of, image_shapes, weights, xform_clip,
                     score_thresh, min_size):
    """b200det_roi_heads_decode: fp32 [R, NC] logits, [R, NC*4] regression, [R, 4] proposals, int32 [R] image index
    -> (boxes [R*(NC-1), 4], scores [R*(NC-1)], valid uint8 [R*(NC-1)])."""
    class_logits, box_regression, proposals = class_logits.contiguous(), box_regression.contiguous(), proposals.contiguous()
    R, NC = class_logits.shape
    dev = class_logits.device
    sizes = torch.tensor([[float(h), float(w)] for h, w in image_shapes], dtype=torch.float32, device=dev)
    n = R * max(NC - 1, 0)
    boxes = torch.empty(n, 4, dtype=torch.float32, device=dev)
    scores = torch.empty(n, dtype=torch.float32, device=dev)
    valid = torch.empty(max(n, 1), dtype=torch.uint8, device=dev)
    w = (C.c_double * 4)(*[float(x) for x in weights])
    with torch.cuda.device(dev):
        _check(lib().b200det_roi_heads_decode(class_logits.data_ptr(), box_regression.data_ptr(), proposals.data_ptr(),
                                              image_of.contiguous().data_ptr(), sizes.data_ptr(), R, NC, w,
                                              float(xform_clip), float(score_thresh), float(min_size),
                                              boxes.data_ptr(), scores.data_ptr(), valid.data_ptr(),
                                              _stream(class_logits)), "b200det_roi_heads_decode")
    return boxes, scores, valid[:n]


def relayout(t, to_channels_last: bool):
    """[N,C,H,W] tensor, dense in one memory format -> the same values dense in the other (b200det_relayout)."""
    N, Cc, H, W = t.shape
    if to_channels_last:
        src = t.contiguous()
        dst = torch.empty_like(src, memory_format=torch.channels_last)
        rows, cols = Cc, H * W
    else:
        src = t.contiguous(memory_format=torch.channels_last)
        dst = torch.empty(t.shape, dtype=t.dtype, device=t.device)
        rows, cols = H * W, Cc
    with torch.cuda.device(t.device):
        _check(lib().b200det_relayout(src.data_ptr(), dst.data_ptr(), _dt(t), N, rows, cols, _stream(t)),
               "b200det_relayout")
    return dst


def roi_pool_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0):
    """(output, argmax int32) in x's memory format (b200det_roi_pool_forward)."""
    layout, x = _layout_of(x)
    r = _arith(rois, x.dtype)
    N, Cc, H, W = x.shape
    K = r.shape[0]
    out = _empty_like_layout((K, Cc, pooled_h, pooled_w), x, layout)
    am = torch.empty_like(out, dtype=torch.int32)  # same memory format as the output
    with torch.cuda.device(x.device):
        _check(lib().b200det_roi_pool_forward(x.data_ptr(), layout, r.data_ptr(), out.data_ptr(), am.data_ptr(), layout,
                                              _dt(x), N, Cc, H, W, K, pooled_h, pooled_w, float(spatial_scale),
                                              _stream(x)), "b200det_roi_pool_forward")
    return out, am


def roi_pool_backward(grad, rois, argmax, pooled_h, pooled_w, N, Cc, H, W, layout=NCHW):
    """grad_input [N,C,H,W] in `layout`; argmax as returned by roi_pool_forward (its own layout is detected)."""
    r = _arith(rois, grad.dtype)
    am_layout, am = _layout_of(argmax)
    gi = _empty_like_layout((N, Cc, H, W), grad, layout)
    with torch.cuda.device(grad.device):
        _check(lib().b200det_roi_pool_backward(grad.data_ptr(), _strides(grad), am.data_ptr(), am_layout, r.data_ptr(),
                                               gi.data_ptr(), layout, _dt(grad), N, Cc, H, W, r.shape[0], pooled_h,
                                               pooled_w, _stream(grad)), "b200det_roi_pool_backward")
    return gi


def ps_roi_pool_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0):
    """(output [K, C/(PH*PW), PH, PW], channel_mapping int32) in x's memory format."""
    layout, x = _layout_of(x)
    r = _arith(rois, x.dtype)
    N, Cc, H, W = x.shape
    K = r.shape[0]
    out = _empty_like_layout((K, Cc // (pooled_h * pooled_w), pooled_h, pooled_w), x, layout)
    cm = torch.empty_like(out, dtype=torch.int32)
    with torch.cuda.device(x.device):
        _check(lib().b200det_ps_roi_pool_forward(x.data_ptr(), layout, r.data_ptr(), out.data_ptr(), cm.data_ptr(),
                                                 layout, _dt(x), N, Cc, H, W, K, pooled_h, pooled_w,
                                                 float(spatial_scale), _stream(x)), "b200det_ps_roi_pool_forward")
    return out, cm


def ps_roi_pool_backward(grad, rois, mapping, spatial_scale, pooled_h, pooled_w, N, Cc, H, W, layout=NCHW):
    r = _arith(rois, grad.dtype)
    cm_layout, cm = _layout_of(mapping)
    gi = _empty_like_layout((N, Cc, H, W), grad, layout)
    with torch.cuda.device(grad.device):
        _check(lib().b200det_ps_roi_pool_backward(grad.data_ptr(), _strides(grad), cm.data_ptr(), cm_layout,
                                                  r.data_ptr(), gi.data_ptr(), layout, _dt(grad), N, Cc, H, W,
                                                  r.shape[0], pooled_h, pooled_w, float(spatial_scale), _stream(grad)),
               "b200det_ps_roi_pool_backward")
    return gi


def ps_roi_align_forward(x, rois, pooled_h, pooled_w, spatial_scale=1.0, sampling_ratio=-1):
    """(output [K, C/(PH*PW), PH, PW], channel_mapping int32) in x's memory format."""
    layout, x = _layout_of(x)
    r = _arith(rois, x.dtype)
    N, Cc, H, W = x.shape
    K = r.shape[0]
    out = _empty_like_layout((K, Cc // (pooled_h * pooled_w), pooled_h, pooled_w), x, layout)
    cm = torch.empty_like(out, dtype=torch.int32)
    with torch.cuda.device(x.device):
        _check(lib().b200det_ps_roi_align_forward(x.data_ptr(), layout, r.data_ptr(), out.data_ptr(), cm.data_ptr(),
                                                  layout, _dt(x), N, Cc, H, W, K, pooled_h, pooled_w,
                                                  float(spatial_scale), int(sampling_ratio), _stream(x)),
               "b200det_ps_roi_align_forward")
    return out, cm


def ps_roi_align_backward(grad, rois, mapping, spatial_scale, pooled_h, pooled_w, sampling_ratio, N, Cc, H, W,
                          layout=NCHW):
    r = _arith(rois, grad.dtype)
    cm_layout, cm = _layout_of(mapping)
    gi = _empty_like_layout((N, Cc, H, W), grad, layout)
    with torch.cuda.device(grad.device):
        _check(lib().b200det_ps_roi_align_backward(grad.data_ptr(), _strides(grad), cm.data_ptr(), cm_layout,
                                                   r.data_ptr(), gi.data_ptr(), layout, _dt(grad), N, Cc, H, W,
                                                   r.shape[0], pooled_h, pooled_w, float(spatial_scale),
                                                   int(sampling_ratio), _stream(grad)),
               "b200det_ps_roi_align_backward")
    return gi


# --------------------------------------------------------------------------- host-buffer entry points
def roi_align_forward_host(x, rois, pooled_h, pooled_w, spatial_scale=1.0, sampling_ratio=-1, aligned=False,
                           device=0, out=None):
    assert not x.is_cuda and not rois.is_cuda
    layout, x = _layout_of(x)
    r = _arith(rois, x.dtype)
    N, Cc, H, W = x.shape
    if out is None:
        out = _empty_like_layout((r.shape[0], Cc, pooled_h, pooled_w), x, layout)
    _check(lib().b200det_roi_align_forward_host(x.data_ptr(), layout, r.data_ptr(), out.data_ptr(), layout, _dt(x), N,
                                                Cc, H, W, r.shape[0], pooled_h, pooled_w, float(spatial_scale),
                                                int(sampling_ratio), int(bool(aligned)), int(device)),
           "b200det_roi_align_forward_host")
    return out


def roi_align_backward_host(grad, rois, spatial_scale, pooled_h, pooled_w, N, Cc, H, W, sampling_ratio, aligned,
                            layout=NCHW, device=0, out=None):
    assert not grad.is_cuda
    r = _arith(rois, grad.dtype)
    gi = out if out is not None else _empty_like_layout((N, Cc, H, W), grad, layout)
    span = 1 + sum((s - 1) * st for s, st in zip(grad.shape, grad.stride())) if grad.numel() else 0
    _check(lib().b200det_roi_align_backward_host(grad.data_ptr(), _strides(grad), span, r.data_ptr(), gi.data_ptr(),
                                                 layout, _dt(grad), N, Cc, H, W, r.shape[0], pooled_h, pooled_w,
                                                 float(spatial_scale), int(sampling_ratio), int(bool(aligned)),
                                                 int(device)),
           "b200det_roi_align_backward_host")
    return gi


def _span(t):
    return 1 + sum((s - 1) * st for s, st in zip(t.shape, t.stride())) if t.numel() else 0


def msroi_align_forward_host(feats, rois, scales, pooled_h, pooled_w, sampling_ratio, aligned=False, k_min=2, k_max=5,
                             canonical_scale=224, canonical_level=4, eps=1e-6, device=0, out=None, levels=None):
    """CPU tensors in, CPU tensors out: (output, levels int32)."""
    layout, f0 = _layout_of(feats[0])
    mf = torch.channels_last if layout == NHWC else torch.contiguous_format
    feats = [f.contiguous(memory_format=mf) for f in feats]
    r = _arith(rois, f0.dtype)
    N, Cc = f0.shape[:2]
    K = r.shape[0]
    if out is None:
        out = _empty_like_layout((K, Cc, pooled_h, pooled_w), f0, layout)
    if levels is None:
        levels = torch.empty(K, dtype=torch.int32)
    ptrs, hs, ws, sc = _level_arrays(feats, scales)
    _check(lib().b200det_msroi_align_forward_host(ptrs, hs, ws, sc, len(feats), layout, r.data_ptr(),
                                                  levels.data_ptr(), out.data_ptr(), layout, _dt(f0), N, Cc, K,
                                                  pooled_h, pooled_w, int(sampling_ratio), int(bool(aligned)), k_min,
                                                  k_max, canonical_scale, canonical_level, float(eps), int(device)),
           "b200det_msroi_align_forward_host")
    return out, levels


def msroi_align_backward_host(grad, rois, levels, scales, shapes, pooled_h, pooled_w, sampling_ratio, aligned=False,
                              layout=NCHW, device=0, outs=None):
    r = _arith(rois, grad.dtype)
    if outs is None:
        outs = [_empty_like_layout(tuple(s), grad, layout) for s in shapes]
    ptrs, hs, ws, sc = _level_arrays(outs, scales)
    N, Cc = shapes[0][:2]
    lv = levels.to(torch.int32).contiguous()
    _check(lib().b200det_msroi_align_backward_host(grad.data_ptr(), _strides(grad), _span(grad), r.data_ptr(),
                                                   lv.data_ptr(), ptrs, hs, ws, sc, len(outs), layout, _dt(grad), N,
                                                   Cc, r.shape[0], pooled_h, pooled_w, int(sampling_ratio),
                                                   int(bool(aligned)), int(device)),
           "b200det_msroi_align_backward_host")
    return outs


def nms_host(boxes, scores, iou_threshold, device=0):
    boxes = boxes.contiguous()
    scores = scores.to(boxes.dtype).contiguous()
    n = boxes.shape[0]
    keep = torch.empty(max(n, 1), dtype=torch.int64)
    cnt = C.c_int64(0)
    _check(lib().b200det_nms_host(boxes.data_ptr(), scores.data_ptr(), _dt(boxes), n, float(iou_threshold),
                                  keep.data_ptr(), C.addressof(cnt), int(device)), "b200det_nms_host")
    return keep[: cnt.value]


def batched_nms_host(boxes, scores, idxs, iou_threshold, mode=BNMS_AUTO, device=0, keep=None):
    boxes = boxes.contiguous()
    scores = scores.to(boxes.dtype).contiguous()
    idxs = idxs.to(torch.int64).contiguous()
    n = boxes.shape[0]
    if keep is None:
        keep = torch.empty(max(n, 1), dtype=torch.int64)
    cnt = C.c_int64(0)
    _check(lib().b200det_batched_nms_host(boxes.data_ptr(), scores.data_ptr(), idxs.data_ptr(), _dt(boxes), n,
                                          float(iou_threshold), int(mode), keep.data_ptr(), C.addressof(cnt),
                                          int(device)), "b200det_batched_nms_host")
    return keep[: cnt.value]


def host_release():
    lib().b200det_host_release()
