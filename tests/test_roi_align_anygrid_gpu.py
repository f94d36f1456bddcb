"""roi_align_anygrid_sm100.cu: channels_last maps with adaptive sampling (sampling_ratio <= 0, the default of
torchvision.ops.roi_align and of every forward in the reference's test/test_ops.py), sampling_ratio > 2, wide
outputs and channel counts that are a multiple of 4 but not of 128.  Checked against the oracle through the C-ABI
with the tolerances of tests/test_roi_align_gpu.py; B200DET_ANYGRID=0 (read per call) sends the same calls to the
generic kernels, which must agree too."""
import pytest
import torch

import oracle
from test_roi_align_gpu import close_bwd, close_fwd, rand_rois

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def capi():
    from cpu_vision_b200 import capi as c

    return c


def cl(t):
    return t.to(DEV).contiguous(memory_format=torch.channels_last)


def kernel_names(fn):
    """Names of the CUDA kernels `fn` launches (None when the profiler has no CUDA activity on this box)."""
    from torch.profiler import ProfilerActivity, profile

    try:
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
        names = [e.key for e in prof.key_averages()]
    except Exception:  # noqa: BLE001
        return None
    return names or None


CASES = [
    # (N, C, H, W, K, PH, PW, scale, sr, aligned)
    (2, 256, 50, 68, 64, 7, 7, 0.0625, -1, True),     # box head, adaptive
    (2, 256, 50, 68, 48, 14, 14, 0.0625, 0, False),   # mask head, adaptive
    (2, 256, 40, 56, 40, 7, 7, 0.25, 3, False),       # sampling_ratio 3
    (1, 128, 30, 30, 24, 4, 6, 0.25, 4, True),        # sampling_ratio 4, PH != PW
    (2, 96, 33, 47, 40, 7, 7, 0.125, -1, False),      # C % 128 != 0: lanes 24..31 idle
    (2, 36, 20, 28, 30, 5, 3, 0.5, 2, True),          # sampling_ratio 2 that the C % 128 kernels do not take
    (1, 160, 25, 25, 20, 7, 7, 0.25, 1, False),       # two channel tiles, the second partly filled
    (1, 32, 64, 64, 16, 40, 33, 1.0, 2, False),       # pooled_width * sampling_ratio > 32, bins smaller than a pixel
    (3, 128, 25, 34, 50, 7, 7, 0.03125, -1, False),   # coarsest level: RoIs smaller than a pixel, grid 1
    (2, 512, 16, 16, 12, 2, 2, 0.5, -1, True),        # few big bins: grids of 4..8
]


@pytest.mark.parametrize("case", CASES)
def test_forward_backward_fp32(capi, case, monkeypatch):
    N, C, H, W, K, PH, PW, scale, sr, aligned = case
    g = torch.Generator().manual_seed(abs(hash(case)) % (2 ** 31))
    x = torch.randn(N, C, H, W, generator=g)
    rois = rand_rois(g, K, N, W / scale, H / scale, lo=1.0 / scale)
    go = torch.randn(K, C, PH, PW, generator=g)
    want = oracle.roi_align_forward(x, rois, PH, PW, scale, sr, aligned)
    want64 = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PW, N, C, H, W, sr, aligned)
    for switch in ("1", "0"):
        monkeypatch.setenv("B200DET_ANYGRID", switch)
        out = capi.roi_align_forward(cl(x), rois.to(DEV), PH, PW, scale, sr, aligned)
        close_fwd(out, want)
        gi = capi.roi_align_backward(cl(go), rois.to(DEV), scale, PH, PW, N, C, H, W, sr, aligned, layout=capi.NHWC)
        close_bwd(gi, want64)


def test_the_anygrid_kernels_are_the_ones_that_run(capi, monkeypatch):
    g = torch.Generator().manual_seed(5)
    x = cl(torch.randn(2, 256, 50, 68, generator=g))
    rois = rand_rois(g, 64, 2, 68 * 16.0, 50 * 16.0, lo=16.0).to(DEV)
    go = cl(torch.randn(64, 256, 7, 7, generator=g))

    def run():
        capi.roi_align_forward(x, rois, 7, 7, 0.0625, -1, True)
        capi.roi_align_backward(go, rois, 0.0625, 7, 7, 2, 256, 50, 68, -1, True, layout=capi.NHWC)

    run()
    names = kernel_names(run)
    if names is None:
        pytest.skip("no CUDA activity from torch.profiler on this box")
    assert any("roi_align_fwd_anygrid_kernel" in n for n in names), names
    assert any("roi_align_bwd_anygrid_kernel" in n for n in names), names
    assert not any("generic" in n for n in names), names
    monkeypatch.setenv("B200DET_ANYGRID", "0")
    names = kernel_names(run)
    assert any("roi_align_fwd_generic_kernel" in n for n in names), names
    assert not any("anygrid" in n for n in names), names


@pytest.mark.parametrize("dtype,rtol,atol,btol", [(torch.bfloat16, 1e-2, 1e-2, 5e-2), (torch.float16, 0.0, 4e-3, 2e-2)])
@pytest.mark.parametrize("C", [256, 40, 36])   # 8 channels per lane (full / partial tile), 4 channels per lane
def test_forward_backward_16bit(capi, dtype, rtol, atol, btol, C):
    g = torch.Generator().manual_seed(C)
    N, H, W, K, PH, scale = 2, 40, 56, 48, 7, 0.125
    x = torch.randn(N, C, H, W, generator=g).to(dtype)
    rois = rand_rois(g, K, N, W / scale, H / scale, lo=1.0 / scale)
    for sr in (-1, 3):
        want = oracle.roi_align_forward(x.float(), rois, PH, PH, scale, sr, True).to(dtype).float()
        out = capi.roi_align_forward(cl(x), rois.to(DEV), PH, PH, scale, sr, True)
        assert out.dtype == dtype
        torch.testing.assert_close(out.cpu().float().contiguous(), want, rtol=rtol, atol=atol)
        go = torch.randn(K, C, PH, PH, generator=g).to(dtype)
        want64 = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PH, N, C, H, W, sr, True)
        gi = capi.roi_align_backward(cl(go), rois.to(DEV), scale, PH, PH, N, C, H, W, sr, True, layout=capi.NHWC)
        assert gi.dtype == dtype
        close_bwd(gi, want64, rtol=btol)


def test_rois_whose_merged_taps_do_not_fit_the_tables(capi):
    """An 800-row map at scale 1: a RoI over the whole height has a 115-sample adaptive grid per bin and more than
    768 merged rows, so its CTA walks the samples instead; its neighbours in the same launch use the tables."""
    g = torch.Generator().manual_seed(11)
    N, C, H, W = 1, 32, 800, 48
    x = torch.randn(N, C, H, W, generator=g)
    rois = torch.tensor([
        [0, 1.0, 2.0, 40.0, 799.0],     # y tables overflow
        [0, 3.0, 100.0, 30.0, 180.0],   # fits
        [0, 0.0, 0.0, 48.0, 800.0],     # the whole map
        [0, 10.0, 790.0, 20.0, 900.0],  # leaves the map at the bottom
    ])
    xw = x.transpose(2, 3).contiguous()           # the same thing along x: a 800-column map
    rois_w = rois[:, [0, 2, 1, 4, 3]].contiguous()
    for xx, rr, hh, ww in ((x, rois, H, W), (xw, rois_w, W, H)):
        for aligned in (False, True):
            want = oracle.roi_align_forward(xx, rr, 7, 7, 1.0, -1, aligned)
            out = capi.roi_align_forward(cl(xx), rr.to(DEV), 7, 7, 1.0, -1, aligned)
            close_fwd(out, want, scale=xx.abs().max().item())
            go = torch.randn(4, C, 7, 7, generator=g)
            want64 = oracle.roi_align_backward(go.double(), rr.double(), 1.0, 7, 7, N, C, hh, ww, -1, aligned)
            gi = capi.roi_align_backward(cl(go), rr.to(DEV), 1.0, 7, 7, N, C, hh, ww, -1, aligned, layout=capi.NHWC)
            close_bwd(gi, want64)


def test_edge_rois_adaptive(capi):
    """Malformed, outside, zero-size, sub-pixel RoIs with adaptive grids (grid 0 or 1, empty samples)."""
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 64, 16, 20, generator=g)
    rois = torch.tensor([
        [0, 10, 12, 4, 3], [1, -50, -50, -20, -20], [0, 100, 100, 150, 150], [1, -3, -3, 6, 6], [0, 7, 7, 7, 7],
        [1, 5.2, 5.3, 5.4, 5.6], [0, 0, 0, 20, 16], [1, 18.5, 14.5, 25, 19], [0, -8, 3, 30, 9],
    ], dtype=torch.float32)
    for sr in (-1, 0, 5):
        for aligned in (True, False):
            want = oracle.roi_align_forward(x, rois, 3, 5, 1.0, sr, aligned)
            out = capi.roi_align_forward(cl(x), rois.to(DEV), 3, 5, 1.0, sr, aligned)
            close_fwd(out, want, scale=x.abs().max().item())
            go = torch.randn(9, 64, 3, 5, generator=g)
            want64 = oracle.roi_align_backward(go.double(), rois.double(), 1.0, 3, 5, 2, 64, 16, 20, sr, aligned)
            gi = capi.roi_align_backward(cl(go), rois.to(DEV), 1.0, 3, 5, 2, 64, 16, 20, sr, aligned, layout=capi.NHWC)
            close_bwd(gi, want64)


def test_backward_accumulates_and_honours_grad_strides(capi):
    g = torch.Generator().manual_seed(4)
    N, C, H, W, K = 2, 128, 24, 24, 20
    rois = rand_rois(g, K, N, W * 4.0, H * 4.0, lo=4.0)
    big = torch.randn(K, 2 * C, 7, 9, generator=g)
    go = cl(big)[:, :C, :, 1:8]                     # channel-contiguous view with padded strides
    want64 = oracle.roi_align_backward(big[:, :C, :, 1:8].double().contiguous(), rois.double(), 0.25, 7, 7, N, C, H, W, -1,
                                       False)
    gi = capi.roi_align_backward(go, rois.to(DEV), 0.25, 7, 7, N, C, H, W, -1, False, layout=capi.NHWC)
    close_bwd(gi, want64)
    seed = cl(torch.randn(N, C, H, W, generator=g))
    acc = capi.roi_align_backward(go, rois.to(DEV), 0.25, 7, 7, N, C, H, W, -1, False, layout=capi.NHWC,
                                  out=seed.clone(memory_format=torch.preserve_format), zero_init=False)
    close_bwd(acc, want64 + seed.cpu().double())
    # a gradient that is not channel-contiguous goes to the generic kernel and still matches
    go_nchw = big[:, :C, :, 1:8].to(DEV)
    gi2 = capi.roi_align_backward(go_nchw, rois.to(DEV), 0.25, 7, 7, N, C, H, W, -1, False, layout=capi.NHWC)
    close_bwd(gi2, want64)


def test_pyramid_with_adaptive_sampling(capi):
    g = torch.Generator().manual_seed(6)
    N, C = 2, 256
    shapes, scales = [(48, 64), (24, 32), (12, 16)], [0.25, 0.125, 0.0625]
    feats = [torch.randn(N, C, h, w, generator=g) for h, w in shapes]
    K = 90
    rois = rand_rois(g, K, N, 256.0, 192.0, lo=8.0)
    levels = oracle.fpn_level_map(rois, 2, 4)
    want = oracle.msroi_align_forward(feats, rois, levels, scales, 7, 7, -1, False)
    out, lv = capi.msroi_align_forward([cl(f) for f in feats], rois.to(DEV), scales, 7, 7, -1, False, k_min=2, k_max=4)
    assert torch.equal(lv.cpu().long(), levels)
    close_fwd(out, want)
    go = torch.randn(K, C, 7, 7, generator=g)
    gis = capi.msroi_align_backward(cl(go), rois.to(DEV), lv, scales, [(N, C, h, w) for h, w in shapes], 7, 7, -1, False,
                                    layout=capi.NHWC)
    for l, (h, w) in enumerate(shapes):
        m = levels == l
        want64 = oracle.roi_align_backward(go[m].double(), rois[m].double(), scales[l], 7, 7, N, C, h, w, -1, False)
        close_bwd(gis[l], want64)


def test_nchw_caller_with_the_default_sampling_ratio_through_the_dispatcher():
    """torchvision.ops.roi_align(x, boxes, 7) -- sampling_ratio -1, NCHW tensors: re-layout + anygrid kernels, NCHW
    results and gradients, equal to the generic strided kernels' within the fp32 tolerances."""
    import cpu_vision_b200  # noqa: F401

    g = torch.Generator().manual_seed(8)
    N, C, H, W, K = 2, 256, 50, 68, 300
    x = torch.randn(N, C, H, W, generator=g)
    rois = rand_rois(g, K, N, W * 8.0, H * 8.0, lo=8.0)
    want = oracle.roi_align_forward(x, rois, 7, 7, 0.125, -1, False)
    xd = x.to(DEV).requires_grad_(True)
    out = torch.ops.torchvision.roi_align(xd, rois.to(DEV), 0.125, 7, 7, -1, False)
    assert out.is_contiguous()
    close_fwd(out.detach(), want)
    go = torch.randn(K, C, 7, 7, generator=g)
    out.backward(go.to(DEV))
    want64 = oracle.roi_align_backward(go.double(), rois.double(), 0.125, 7, 7, N, C, H, W, -1, False)
    assert xd.grad.is_contiguous()
    close_bwd(xd.grad, want64)
