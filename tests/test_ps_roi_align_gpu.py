"""torchvision::ps_roi_align on the GPU against the oracle and golden outputs of the
reference's own ops.ps_roi_align, through the C-ABI and the dispatcher.  Bar: forward rtol 1e-5 (fp32; the GPU
may contract w*v products into FMAs), channel_mapping integer-exact, backward within the atomic-reordering
tolerance (rtol 1e-5 fp32), NaN pattern of empty adaptive grids identical."""
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def capi():
    from cpu_vision_b200 import capi as c

    return c


def _mf(t, channels_last):
    return t.contiguous(memory_format=torch.channels_last) if channels_last else t.contiguous()


def _close(got, want, rtol):
    assert torch.equal(got.isnan(), want.isnan())
    g, w = got.nan_to_num().double(), want.nan_to_num().double()
    assert (g - w).abs().max().item() <= rtol * max(w.abs().max().item(), 1.0)


@pytest.mark.parametrize("channels_last", [False, True])
def test_golden(golden, capi, channels_last):
    d = golden("ps_roi_align.pt")
    rois = d["rois"]
    for c in d["cases"]:
        x = c["x"]
        out, cm = capi.ps_roi_align_forward(_mf(x.to(DEV), channels_last), rois.to(DEV), c["ph"], c["pw"], c["scale"],
                                            c["sr"])
        _close(out.cpu().contiguous(), c["out"], 1e-5)
        _, want_cm = oracle.ps_roi_align_forward(x, rois, c["ph"], c["pw"], c["scale"], c["sr"])
        assert torch.equal(cm.cpu().contiguous(), want_cm)
        go = c["grad_out"]
        gi = capi.ps_roi_align_backward(_mf(go.to(DEV), channels_last), rois.to(DEV), cm, c["scale"], c["ph"], c["pw"],
                                        c["sr"], *x.shape, layout=capi.NHWC if channels_last else capi.NCHW)
        _close(gi.cpu().contiguous(), c["grad_in"], 1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.float16, torch.bfloat16])
def test_random_vs_oracle(capi, dtype):
    g = torch.Generator().manual_seed(43)
    N, H, W, K = 3, 37, 45, 150
    rois = torch.rand(K, 5, generator=g) * 160 - 10
    rois[:, 0] = torch.randint(0, N, (K,), generator=g).float()
    rois[:, 3:] = rois[:, 1:3] + torch.rand(K, 2, generator=g) * 120 + 1
    ref_dtype = torch.float64 if dtype == torch.float64 else torch.float32
    ftol = {torch.float32: 1e-5, torch.float64: 1e-12, torch.float16: 4e-3, torch.bfloat16: 2e-2}[dtype]
    btol = {torch.float32: 1e-5, torch.float64: 1e-12, torch.float16: 2e-2, torch.bfloat16: 8e-2}[dtype]
    for ph, pw, mul, scale, sr in ((7, 7, 2, 0.25, 2), (2, 3, 10, 0.5, -1), (3, 3, 21, 0.125, 3)):
        C = mul * ph * pw
        x = torch.randn(N, C, H, W, generator=g).to(dtype)
        want, want_cm = oracle.ps_roi_align_forward(x.to(ref_dtype), rois.to(ref_dtype), ph, pw, scale, sr)
        out, cm = capi.ps_roi_align_forward(x.to(DEV), rois.to(DEV), ph, pw, scale, sr)
        assert torch.equal(cm.cpu(), want_cm)
        _close(out.cpu().to(ref_dtype), want, ftol)
        go = torch.randn(K, mul, ph, pw, generator=g).to(dtype)
        gi = capi.ps_roi_align_backward(go.to(DEV), rois.to(DEV), cm, scale, ph, pw, sr, N, C, H, W)
        want_gi = oracle.ps_roi_align_backward(go.double(), rois.double(), want_cm, scale, ph, pw, sr, N, C, H, W)
        _close(gi.cpu().double(), want_gi, btol)


@pytest.mark.parametrize("channels_last", [False, True])
def test_dispatcher_module_and_autograd(golden, channels_last):
    from cpu_vision_b200 import PSRoIAlign

    d = golden("ps_roi_align.pt")
    c = d["cases"][0]
    xd = _mf(c["x"].to(DEV), channels_last).requires_grad_(True)
    m = PSRoIAlign((c["ph"], c["pw"]), c["scale"], c["sr"])
    assert "sampling_ratio=2" in repr(m)
    y = m(xd, d["rois"].to(DEV))
    _close(y.detach().cpu().contiguous(), c["out"], 1e-5)
    y.backward(c["grad_out"].to(DEV))
    _close(xd.grad.cpu().contiguous(), c["grad_in"], 1e-5)
    with pytest.raises(RuntimeError, match="multiple of pooling height"):
        torch.ops.torchvision.ps_roi_align(xd.detach()[:, :5], d["rois"].to(DEV), 1.0, 3, 3, 2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(c["x"], d["rois"])
