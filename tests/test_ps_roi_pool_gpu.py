"""torchvision::ps_roi_pool on the GPU against the oracle and golden outputs of the
reference's own ops.ps_roi_pool, through the C-ABI.  Bar: forward bit-exact in fp32 / fp64 (a thread sums
its bin in the reference's h-major order), channel_mapping integer-exact, backward within the
atomic-reordering tolerance (rtol 1e-5 fp32)."""
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


@pytest.mark.parametrize("channels_last", [False, True])
def test_golden(golden, capi, channels_last):
    d = golden("ps_roi_pool.pt")
    rois = d["rois"]
    for c in d["cases"]:
        x = c["x"]
        out, cm = capi.ps_roi_pool_forward(_mf(x.to(DEV), channels_last), rois.to(DEV), c["ph"], c["pw"], c["scale"])
        assert torch.equal(out.cpu().contiguous(), c["out"]), (c["ph"], c["scale"])
        _, want_cm = oracle.ps_roi_pool_forward(x, rois, c["ph"], c["pw"], c["scale"])
        assert torch.equal(cm.cpu().contiguous(), want_cm)
        gi = capi.ps_roi_pool_backward(_mf(c["grad_out"].to(DEV), channels_last), rois.to(DEV), cm, c["scale"], c["ph"],
                                       c["pw"], *x.shape, layout=capi.NHWC if channels_last else capi.NCHW)
        torch.testing.assert_close(gi.cpu().contiguous(), c["grad_in"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.float16, torch.bfloat16])
def test_random_vs_oracle(capi, dtype):
    g = torch.Generator().manual_seed(41)
    N, H, W, K = 3, 37, 45, 150
    rois = torch.rand(K, 5, generator=g) * 160 - 10
    rois[:, 0] = torch.randint(0, N, (K,), generator=g).float()
    rois[:, 3:] = rois[:, 1:3] + torch.rand(K, 2, generator=g) * 120 - 5
    ref_dtype = torch.float64 if dtype == torch.float64 else torch.float32
    for ph, pw, mul, scale in ((7, 7, 2, 0.25), (2, 3, 10, 0.5), (3, 3, 21, 0.125)):
        C = mul * ph * pw
        x = torch.randn(N, C, H, W, generator=g).to(dtype)
        want, want_cm = oracle.ps_roi_pool_forward(x.to(ref_dtype), rois.to(ref_dtype), ph, pw, scale)
        out, cm = capi.ps_roi_pool_forward(x.to(DEV), rois.to(DEV), ph, pw, scale)
        assert torch.equal(cm.cpu(), want_cm)
        if dtype in (torch.float32, torch.float64):
            assert torch.equal(out.cpu(), want)
        else:
            torch.testing.assert_close(out.cpu().float(), want, rtol=1e-2, atol=1e-2)
        go = torch.randn(K, mul, ph, pw, generator=g).to(dtype)
        gi = capi.ps_roi_pool_backward(go.to(DEV), rois.to(DEV), cm, scale, ph, pw, N, C, H, W)
        want_gi = oracle.ps_roi_pool_backward(go.double(), rois.double(), want_cm, scale, ph, pw, N, C, H, W)
        tol = {torch.float32: 1e-5, torch.float64: 1e-12, torch.float16: 2e-2, torch.bfloat16: 8e-2}[dtype]
        err = (gi.cpu().double() - want_gi).abs().max().item()
        assert err <= tol * max(want_gi.abs().max().item(), 1.0), (dtype, err)


def test_errors_and_empty(capi):
    x = torch.randn(1, 10, 8, 8, device=DEV)
    rois = torch.tensor([[0, 0.0, 0.0, 6.0, 6.0]], device=DEV)
    with pytest.raises(capi.B200DetError, match="multiple of pooling height"):
        capi.ps_roi_pool_forward(x, rois, 3, 3, 1.0)
    out, cm = capi.ps_roi_pool_forward(x[:, :9], rois[:0], 3, 3, 1.0)
    assert out.shape == (0, 1, 3, 3)


@pytest.mark.parametrize("channels_last", [False, True])
def test_dispatcher_module_and_autograd(golden, channels_last):
    from cpu_vision_b200 import PSRoIPool

    d = golden("ps_roi_pool.pt")
    c = d["cases"][0]
    xd = _mf(c["x"].to(DEV), channels_last).requires_grad_(True)
    m = PSRoIPool((c["ph"], c["pw"]), c["scale"])
    y = m(xd, d["rois"].to(DEV))
    assert torch.equal(y.detach().cpu().contiguous(), c["out"])
    y.backward(c["grad_out"].to(DEV))
    torch.testing.assert_close(xd.grad.cpu().contiguous(), c["grad_in"], rtol=1e-5, atol=1e-5)
    with pytest.raises(RuntimeError, match="multiple of pooling height"):
        torch.ops.torchvision.ps_roi_pool(xd.detach()[:, :5], d["rois"].to(DEV), 1.0, 3, 3)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(c["x"], d["rois"])
