"""Autocast (rows R7 / N7 of SURVEY section 8a): mirrors test/test_ops.py:493-508 (`test_autocast`: for x and rois
in {float, half}^2 the output dtype equals x's) and the reference's semantics
roi_align_autocast = ref_fp32(x.float(), rois.float()).to(x.dtype) (csrc/ops/autocast/roi_align_kernel.cpp:12-31),
computed here WITHOUT the fp32 copy of the feature map: 16-bit features reach the native kernels."""
import pytest
import torch

import oracle
from test_roi_align_gpu import DEV, rand_rois

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("amp_dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("x_dtype", ["float", "low"])
@pytest.mark.parametrize("rois_dtype", ["float", "low"])
@pytest.mark.parametrize("channels_last", [False, True])
def test_roi_align_autocast(amp_dtype, x_dtype, rois_dtype, channels_last):
    import cpu_vision_b200  # noqa: F401

    g = torch.Generator().manual_seed(41)
    N, C, H, W, K, PH = 2, 256, 40, 56, 120, 7
    xd = amp_dtype if x_dtype == "low" else torch.float32
    rd = torch.float16 if rois_dtype == "low" else torch.float32   # the reference's test uses half rois
    x = torch.randn(N, C, H, W, generator=g).to(xd)
    rois = rand_rois(g, K, N, W * 4.0, H * 4.0, lo=8.0).to(rd)
    want = oracle.roi_align_forward(x.float(), rois.float(), PH, PH, 0.25, 2, False).to(xd)
    xg = x.to(DEV)
    if channels_last:
        xg = xg.contiguous(memory_format=torch.channels_last)
    xg.requires_grad_(True)
    with torch.autocast("cuda", dtype=amp_dtype):
        y = torch.ops.torchvision.roi_align(xg, rois.to(DEV), 0.25, PH, PH, 2, False)
    assert y.dtype == xd
    tol = dict(rtol=1e-5, atol=1e-5) if xd == torch.float32 else dict(rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(y.detach().float().cpu().contiguous(), want.float(), **tol)
    go = torch.randn(K, C, PH, PH, generator=g).to(xd)
    y.backward(go.to(DEV))
    assert xg.grad is not None and xg.grad.dtype == xd
    want_g = oracle.roi_align_backward(go.double(), rois.double(), 0.25, PH, PH, N, C, H, W, 2, False)
    err = (xg.grad.double().cpu() - want_g).abs().max().item()
    assert err <= (1e-4 if xd == torch.float32 else 5e-2) * want_g.abs().max().item()


@pytest.mark.parametrize("amp_dtype", [torch.float16, torch.bfloat16])
def test_nms_autocast(amp_dtype):
    """test_ops.py:856-861 (`TestNMS.test_autocast`): nms of half boxes under autocast equals nms of the widened
    boxes; mixed dtypes are cast to fp32 like the reference's wrapper does."""
    import cpu_vision_b200  # noqa: F401

    g = torch.Generator().manual_seed(42)
    b = torch.rand(800, 4, generator=g) * 100
    b[:, 2:] += b[:, :2]
    s = torch.randperm(800, generator=g).float() / 800
    for bd, sd in ((amp_dtype, amp_dtype), (amp_dtype, torch.float32), (torch.float32, torch.float32)):
        bb, ss = b.to(bd), s.to(sd)
        want = oracle.nms(bb.float(), ss.float(), 0.5)
        with torch.autocast("cuda", dtype=amp_dtype):
            got = torch.ops.torchvision.nms(bb.to(DEV), ss.to(DEV), 0.5)
        assert torch.equal(got.cpu(), want), (bd, sd)


def test_autocast_msroi_module_keeps_16bit_features():
    """MultiScaleRoIAlign under autocast with bf16 FPN features (what an AMP backbone hands over): the pooled
    output is bf16 and matches the reference's autocast semantics."""
    import cpu_vision_b200 as m
    from collections import OrderedDict

    g = torch.Generator().manual_seed(43)
    shapes = [(50, 68), (25, 34)]
    feats = OrderedDict((str(i), torch.randn(2, 256, h, w, generator=g).bfloat16().to(DEV)) for i, (h, w) in enumerate(shapes))
    boxes = [rand_rois(g, 50, 1, 272.0, 200.0, lo=8.0)[:, 1:].to(DEV) for _ in range(2)]
    pool = m.MultiScaleRoIAlign(["0", "1"], 7, 2)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = pool(feats, boxes, [(200, 272)] * 2)
    assert out.dtype == torch.bfloat16
    rois = torch.cat([torch.cat([torch.full((50, 1), float(i)), b.cpu()], 1) for i, b in enumerate(boxes)])
    lv = oracle.fpn_level_map(rois, 2, 3)
    want = oracle.msroi_align_forward([f.float().cpu() for f in feats.values()], rois, lv, [0.25, 0.125], 7, 7, 2, False)
    torch.testing.assert_close(out.float().cpu().contiguous(), want.bfloat16().float(), rtol=1e-2, atol=1e-2)
