"""No-GPU checks: the C-ABI library loads, exports every symbol the header
declares, rejects bad arguments before touching the device, and the dispatcher
shim registers the reference's operator names under the CUDA key."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_detection_ops.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200det_\w+)\s*\(", src)))


def test_header_symbols_are_exported():
    from cpu_vision_b200 import _lib, capi

    lib = ctypes.CDLL(_lib.CAPI_PATH)
    syms = declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/b200_detection_ops.h but not exported"
    assert sorted(capi.SYMBOLS) == syms, "capi.py binding list and header disagree"
    assert capi.lib().b200det_version() == 100


def test_argument_errors_without_gpu():
    from cpu_vision_b200 import capi

    l = capi.lib()
    # unknown dtype / negative size / bad layout are rejected before any CUDA call
    assert l.b200det_roi_align_forward(None, 0, None, None, 0, 99, 1, 1, 1, 1, 1, 1, 1, 1.0, 2, 0, None) == -2
    assert b"dtype" in l.b200det_last_error()
    assert l.b200det_roi_align_forward(None, 0, None, None, 0, 0, 1, 1, 1, 1, -1, 1, 1, 1.0, 2, 0, None) == -1
    assert l.b200det_roi_align_forward(None, 7, None, None, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1.0, 2, 0, None) == -1
    assert l.b200det_roi_align_forward(None, 0, None, None, 0, 0, 1, 1, 4, 4, 3, 2, 2, 1.0, 2, 0, None) == -1  # null ptr
    assert l.b200det_nms(None, None, 0, -5, 0.5, None, None, None, 0, None) == -1
    assert l.b200det_batched_nms(None, None, None, 0, 4, 0.5, 5, None, ctypes.c_void_p(8), None, 0, None) == -1
    assert l.b200det_nms_workspace_bytes(5000) > 5000 * 79 * 8
    assert l.b200det_nms_workspace_bytes(0) > 0
    # the round-2 entry points: the push forward refuses more than 7 peers and null peer arrays, the per-level top-k
    # refuses k beyond 2048 (ENOSUP: the caller keeps torch.topk), level ends that do not end at A, a wrong P
    le = (ctypes.c_int32 * 2)(100, 300)
    assert l.b200det_msroi_align_forward_push(None, None, None, None, 1, 1, None, None, 0, None, None, 8, 1, 0, 1, 256,
                                              4, 7, 7, 2, 0, 2, 5, 224, 4, 1e-6, None) == -1
    assert b"at most 7 peer outputs" in l.b200det_last_error()
    assert l.b200det_msroi_align_forward_push(None, None, None, None, 1, 1, None, None, 0, None, None, 2, 1, 0, 1, 256,
                                              4, 7, 7, 2, 0, 2, 5, 224, 4, 1e-6, None) == -1
    assert l.b200det_segmented_topk(None, 2, 300, le, 2, 4096, None, 400, None) == -2
    assert l.b200det_segmented_topk(None, 2, 301, le, 2, 100, None, 200, None) == -1
    assert l.b200det_segmented_topk(None, 2, 300, le, 2, 150, None, 999, None) == -1
    assert b"P must be" in l.b200det_last_error()
    assert l.b200det_segmented_topk(None, 0, 300, le, 2, 150, None, 250, None) == 0      # no images: nothing to do
    assert l.b200det_peer_alloc(0, None) == -1 and l.b200det_peer_export(None, None) == -1


def test_dispatcher_registration():
    import cpu_vision_b200  # noqa: F401  (loads the shim)

    for name in ("torchvision::roi_align", "torchvision::_roi_align_backward", "torchvision::nms",
                 "b200det::msroi_align", "b200det::msroi_align_backward", "b200det::batched_nms"):
        dump = torch._C._dispatch_dump(name)
        assert "CUDA: registered at" in dump and "torch_shim.cpp" in dump, dump
    s = str(torch.ops.torchvision.roi_align.default._schema)
    assert "float spatial_scale, SymInt pooled_height, SymInt pooled_width, int sampling_ratio, bool aligned" in s
    s = str(torch.ops.torchvision._roi_align_backward.default._schema)
    assert "SymInt batch_size, SymInt channels, SymInt height, SymInt width, int sampling_ratio, bool aligned" in s
    assert "Tensor dets, Tensor scores, float iou_threshold" in str(torch.ops.torchvision.nms.default._schema)


def test_no_cpu_fallback():
    """The product refuses CPU tensors instead of silently computing somewhere else."""
    import cpu_vision_b200 as m

    x = torch.randn(1, 4, 8, 8)
    rois = torch.tensor([[0.0, 1, 1, 5, 5]])
    with pytest.raises(RuntimeError):
        m.roi_align(x, rois, 2)
    with pytest.raises(RuntimeError):
        m.nms(rois[:, 1:], torch.tensor([0.5]), 0.5)
    with pytest.raises((RuntimeError, NotImplementedError)):
        torch.ops.torchvision.roi_align(x, rois, 1.0, 2, 2, 2, False)


def test_product_does_not_import_oracle():
    import subprocess
    import sys

    code = ("import sys; import cpu_vision_b200, cpu_vision_b200.capi, cpu_vision_b200.sharding; "
            "bad=[m for m in sys.modules if m=='oracle' or m.startswith('oracle.')]; assert not bad, bad")
    subprocess.check_call([sys.executable, "-c", code], cwd=ROOT)
    for dirpath, _, files in os.walk(os.path.join(ROOT, "cpu_vision_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in txt and "import oracle" not in txt and "libtvref" not in txt, f


def test_fake_kernels_for_compile():
    import cpu_vision_b200  # noqa: F401
    from torch._subclasses.fake_tensor import FakeTensorMode

    with FakeTensorMode():
        x = torch.empty(2, 8, 16, 16, device="cuda")
        r = torch.empty(5, 5, device="cuda")
        y = torch.ops.torchvision.roi_align(x, r, 0.5, 7, 7, 2, True)
        assert tuple(y.shape) == (5, 8, 7, 7)
        out, lv = torch.ops.b200det.msroi_align([x, x], r, [0.5, 0.25], 7, 7, 2, False, 2, 3, 224, 4, 1e-6)
        assert tuple(out.shape) == (5, 8, 7, 7) and lv.dtype == torch.int32
