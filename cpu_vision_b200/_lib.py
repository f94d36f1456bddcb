"""Loads the native libraries and completes the operator registration.

libb200det.so      C-ABI kernels (include/b200_detection_ops.h), no torch dependency
b200det_torch.so   TORCH_LIBRARY_IMPL(torchvision, CUDA, ...) shim over it

There is no fallback: if either library is missing the import fails loudly
(build them with `python -c "import __graft_entry__ as g; g.build()"` or
`make -C cpu_vision_b200/csrc`).

When the shim found no torchvision schema in the dispatcher (no reference build
loaded before us) it defined `torchvision::roi_align`, `_roi_align_backward`
and `nms` itself; in that case the Autograd and Meta companions the reference
ships (csrc/ops/autograd/roi_align_kernel.cpp:111-164,
torchvision/_meta_registrations.py:25-51,163-174) are registered here so the
ops behave the same stand-alone.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
CAPI_PATH = os.path.join(_HERE, "libb200det.so")
SHIM_PATH = os.path.join(_HERE, "b200det_torch.so")

_state = {"capi": None, "shim": None, "defined": 0}


def _missing(path):
    raise ImportError(
        f"{path} not found: the CUDA extension is not built. Run "
        "`python -c 'import __graft_entry__ as g; g.build()'` (or `make -C cpu_vision_b200/csrc`). "
        "cpu_vision_b200 has no CPU or PyTorch fallback."
    )


def capi():
    """ctypes handle of libb200det.so."""
    if _state["capi"] is None:
        if not os.path.exists(CAPI_PATH):
            _missing(CAPI_PATH)
        _state["capi"] = ctypes.CDLL(CAPI_PATH, mode=ctypes.RTLD_GLOBAL)
    return _state["capi"]


def load():
    """Load the dispatcher shim (idempotent)."""
    if _state["shim"] is not None:
        return _state["shim"]
    capi()
    if not os.path.exists(SHIM_PATH):
        _missing(SHIM_PATH)
    torch.ops.load_library(SHIM_PATH)
    shim = ctypes.CDLL(SHIM_PATH)
    shim.b200det_shim_defined_schemas.restype = ctypes.c_int
    _state["defined"] = shim.b200det_shim_defined_schemas()
    _state["shim"] = shim
    _register_companions(_state["defined"])
    return shim


def defined_schemas():
    load()
    return _state["defined"]


def _register_companions(defined):
    lib = torch.library

    # --- fused entry points in our own namespace always need their companions
    @lib.register_fake("b200det::msroi_align")
    def _(feats, rois, scales, pooled_height, pooled_width, sampling_ratio, aligned, k_min, k_max,
          canonical_scale, canonical_level, eps):
        f0 = feats[0]
        out = f0.new_empty((rois.size(0), f0.size(1), pooled_height, pooled_width))
        return out, rois.new_empty((rois.size(0),), dtype=torch.int32)

    @lib.register_fake("b200det::msroi_align_backward")
    def _(grad, rois, levels, scales, heights, widths, batch_size, channels, pooled_height, pooled_width,
          sampling_ratio, aligned, channels_last):
        return [grad.new_empty((batch_size, channels, h, w)) for h, w in zip(heights, widths)]

    @lib.register_fake("b200det::fpn_level_map")
    def _(rois, k_min, k_max, canonical_scale, canonical_level, eps):
        return rois.new_empty((rois.size(0),), dtype=torch.int32)

    @lib.register_fake("b200det::batched_nms")
    def _(boxes, scores, idxs, iou_threshold, mode):
        ctx = torch._custom_ops.get_ctx()
        n = ctx.create_unbacked_symint()
        return boxes.new_empty((n,), dtype=torch.int64)

    def _ms_setup(ctx, inputs, output):
        feats, rois, scales, ph, pw, sr, aligned = inputs[:7]
        ctx.save_for_backward(rois, output[1])
        ctx.meta = (list(scales), [f.shape[2] for f in feats], [f.shape[3] for f in feats],
                    feats[0].shape[0], feats[0].shape[1], ph, pw, sr, aligned,
                    feats[0].dim() == 4 and feats[0].is_contiguous(memory_format=torch.channels_last)
                    and not feats[0].is_contiguous())

    def _ms_backward(ctx, grad_out, _grad_levels):
        rois, levels = ctx.saved_tensors
        scales, hs, ws, n, c, ph, pw, sr, aligned, cl = ctx.meta
        grads = torch.ops.b200det.msroi_align_backward(grad_out, rois, levels, scales, hs, ws, n, c, ph, pw, sr,
                                                        aligned, cl)
        return (list(grads),) + (None,) * 11

    lib.register_autograd("b200det::msroi_align", _ms_backward, setup_context=_ms_setup)

    # --- torchvision:: companions only when nobody else provided them
    if defined & 1:
        @lib.register_fake("torchvision::roi_align")
        def _(input, rois, spatial_scale, pooled_height, pooled_width, sampling_ratio, aligned):
            torch._check(rois.size(1) == 5, lambda: "rois must have shape as Tensor[K, 5]")
            return input.new_empty((rois.size(0), input.size(1), pooled_height, pooled_width))

        @lib.register_fake("torchvision::_roi_align_backward")
        def _(grad, rois, spatial_scale, pooled_height, pooled_width, batch_size, channels, height, width,
              sampling_ratio, aligned):
            return grad.new_empty((batch_size, channels, height, width))

        def _setup(ctx, inputs, output):
            input, rois, scale, ph, pw, sr, aligned = inputs
            ctx.save_for_backward(rois)
            ctx.meta = (scale, ph, pw, tuple(input.shape), sr, aligned)

        def _backward(ctx, grad):
            (rois,) = ctx.saved_tensors
            scale, ph, pw, shape, sr, aligned = ctx.meta
            gi = torch.ops.torchvision._roi_align_backward(grad, rois, scale, ph, pw, shape[0], shape[1],
                                                           shape[2], shape[3], sr, aligned)
            return gi, None, None, None, None, None, None

        lib.register_autograd("torchvision::roi_align", _backward, setup_context=_setup)

        def _no_double_backward(ctx, *grads):
            # autograd/roi_align_kernel.cpp:104-108
            raise RuntimeError("double backwards on roi_align not supported")

        lib.register_autograd("torchvision::_roi_align_backward", _no_double_backward,
                              setup_context=lambda ctx, inputs, output: None)

    if defined & 2:
        @lib.register_fake("torchvision::nms")
        def _(dets, scores, iou_threshold):
            torch._check(dets.dim() == 2, lambda: f"boxes should be a 2d tensor, got {dets.dim()}D")
            torch._check(dets.size(1) == 4,
                         lambda: f"boxes should have 4 elements in dimension 1, got {dets.size(1)}")
            torch._check(scores.dim() == 1, lambda: f"scores should be a 1d tensor, got {scores.dim()}")
            ctx = torch._custom_ops.get_ctx()
            n = ctx.create_unbacked_symint()
            return dets.new_empty((n,), dtype=torch.int64)

    if defined & 4:
        # csrc/ops/autograd/roi_pool_kernel.cpp:11-152, torchvision/_meta_registrations.py (roi_pool)
        @lib.register_fake("torchvision::roi_pool")
        def _(input, rois, spatial_scale, pooled_height, pooled_width):
            torch._check(rois.size(1) == 5, lambda: "rois must have shape as Tensor[K, 5]")
            shape = (rois.size(0), input.size(1), pooled_height, pooled_width)
            return input.new_empty(shape), input.new_empty(shape, dtype=torch.int32)

        @lib.register_fake("torchvision::_roi_pool_backward")
        def _(grad, rois, argmax, spatial_scale, pooled_height, pooled_width, batch_size, channels, height, width):
            return grad.new_empty((batch_size, channels, height, width))

        def _pool_setup(ctx, inputs, output):
            input, rois, scale, ph, pw = inputs
            ctx.save_for_backward(rois, output[1])
            ctx.mark_non_differentiable(output[1])
            ctx.meta = (scale, ph, pw, tuple(input.shape))

        def _pool_backward(ctx, grad, _grad_argmax):
            rois, argmax = ctx.saved_tensors
            scale, ph, pw, shape = ctx.meta
            gi = torch.ops.torchvision._roi_pool_backward(grad, rois, argmax, scale, ph, pw, shape[0], shape[1],
                                                          shape[2], shape[3])
            return gi, None, None, None, None

        lib.register_autograd("torchvision::roi_pool", _pool_backward, setup_context=_pool_setup)

    if defined & 8:
        # csrc/ops/autograd/ps_roi_pool_kernel.cpp, torchvision/_meta_registrations.py (ps_roi_pool)
        @lib.register_fake("torchvision::ps_roi_pool")
        def _(input, rois, spatial_scale, pooled_height, pooled_width):
            torch._check(rois.size(1) == 5, lambda: "rois must have shape as Tensor[K, 5]")
            torch._check(input.size(1) % (pooled_height * pooled_width) == 0,
                         lambda: "input channels must be a multiple of pooling height * pooling width")
            shape = (rois.size(0), input.size(1) // (pooled_height * pooled_width), pooled_height, pooled_width)
            return input.new_empty(shape), input.new_empty(shape, dtype=torch.int32)

        @lib.register_fake("torchvision::_ps_roi_pool_backward")
        def _(grad, rois, channel_mapping, spatial_scale, pooled_height, pooled_width, batch_size, channels, height,
              width):
            return grad.new_empty((batch_size, channels, height, width))

        def _ps_setup(ctx, inputs, output):
            input, rois, scale, ph, pw = inputs
            ctx.save_for_backward(rois, output[1])
            ctx.mark_non_differentiable(output[1])
            ctx.meta = (scale, ph, pw, tuple(input.shape))

        def _ps_backward(ctx, grad, _grad_mapping):
            rois, mapping = ctx.saved_tensors
            scale, ph, pw, shape = ctx.meta
            gi = torch.ops.torchvision._ps_roi_pool_backward(grad, rois, mapping, scale, ph, pw, shape[0], shape[1],
                                                             shape[2], shape[3])
            return gi, None, None, None, None

        lib.register_autograd("torchvision::ps_roi_pool", _ps_backward, setup_context=_ps_setup)

    if defined & 16:
        # csrc/ops/autograd/ps_roi_align_kernel.cpp, torchvision/_meta_registrations.py (ps_roi_align)
        @lib.register_fake("torchvision::ps_roi_align")
        def _(input, rois, spatial_scale, pooled_height, pooled_width, sampling_ratio):
            torch._check(rois.size(1) == 5, lambda: "rois must have shape as Tensor[K, 5]")
            torch._check(input.size(1) % (pooled_height * pooled_width) == 0,
                         lambda: "input channels must be a multiple of pooling height * pooling width")
            shape = (rois.size(0), input.size(1) // (pooled_height * pooled_width), pooled_height, pooled_width)
            return input.new_empty(shape), input.new_empty(shape, dtype=torch.int32)

        @lib.register_fake("torchvision::_ps_roi_align_backward")
        def _(grad, rois, channel_mapping, spatial_scale, pooled_height, pooled_width, sampling_ratio, batch_size,
              channels, height, width):
            return grad.new_empty((batch_size, channels, height, width))

        def _psa_setup(ctx, inputs, output):
            input, rois, scale, ph, pw, sr = inputs
            ctx.save_for_backward(rois, output[1])
            ctx.mark_non_differentiable(output[1])
            ctx.meta = (scale, ph, pw, sr, tuple(input.shape))

        def _psa_backward(ctx, grad, _grad_mapping):
            rois, mapping = ctx.saved_tensors
            scale, ph, pw, sr, shape = ctx.meta
            gi = torch.ops.torchvision._ps_roi_align_backward(grad, rois, mapping, scale, ph, pw, sr, shape[0],
                                                              shape[1], shape[2], shape[3])
            return gi, None, None, None, None, None

        lib.register_autograd("torchvision::ps_roi_align", _psa_backward, setup_context=_psa_setup)
