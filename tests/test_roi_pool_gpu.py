"""torchvision::roi_pool on the GPU against the oracle and golden outputs of the
reference's own ops.roi_pool, through the C-ABI.  Bar: forward values and argmax identical (max pooling
selects, it does not round; the first maximum in row-major order wins in both); backward within the
atomic-reordering tolerance (rtol 1e-5 fp32: a pixel collects at most a few dozen addends)."""
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
    d = golden("roi_pool.pt")
    x, rois = d["x"], d["rois"]
    N, C, H, W = x.shape
    xd = _mf(x.to(DEV), channels_last)
    for c in d["cases"]:
        out, am = capi.roi_pool_forward(xd, rois.to(DEV), c["ph"], c["pw"], c["scale"])
        assert torch.equal(out.cpu().contiguous(), c["out"]), (c["ph"], c["scale"])
        _, want_am = oracle.roi_pool_forward(x, rois, c["ph"], c["pw"], c["scale"])
        assert torch.equal(am.cpu().contiguous(), want_am)
        gi = capi.roi_pool_backward(_mf(c["grad_out"].to(DEV), channels_last), rois.to(DEV), am, c["ph"], c["pw"],
                                    N, C, H, W, layout=capi.NHWC if channels_last else capi.NCHW)
        torch.testing.assert_close(gi.cpu().contiguous(), c["grad_in"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("channels_last", [False, True])
def test_random_vs_oracle(capi, dtype, channels_last):
    g = torch.Generator().manual_seed(31)
    N, C, H, W, K = 3, 70, 37, 45, 200
    x = torch.randn(N, C, H, W, generator=g).to(dtype)
    rois = torch.rand(K, 5, generator=g) * 160 - 10
    rois[:, 0] = torch.randint(0, N, (K,), generator=g).float()
    rois[:, 3:] = rois[:, 1:3] + torch.rand(K, 2, generator=g) * 120 - 5  # some malformed, some outside
    ref_dtype = torch.float64 if dtype == torch.float64 else torch.float32
    for ph, pw, scale in ((7, 7, 0.25), (2, 3, 0.5), (14, 14, 0.125)):
        want, want_am = oracle.roi_pool_forward(x.to(ref_dtype), rois.to(ref_dtype), ph, pw, scale)
        out, am = capi.roi_pool_forward(_mf(x.to(DEV), channels_last), rois.to(DEV), ph, pw, scale)
        assert torch.equal(out.cpu().contiguous().to(ref_dtype), want.to(dtype).to(ref_dtype))
        assert torch.equal(am.cpu().contiguous(), want_am)
        go = torch.randn(K, C, ph, pw, generator=g).to(dtype)
        gi = capi.roi_pool_backward(_mf(go.to(DEV), channels_last), rois.to(DEV), am, ph, pw, N, C, H, W,
                                    layout=capi.NHWC if channels_last else capi.NCHW)
        want_gi = oracle.roi_pool_backward(go.double(), rois.double(), want_am, ph, pw, N, C, H, W)
        tol = {torch.float32: 1e-5, torch.float64: 1e-12, torch.float16: 2e-2, torch.bfloat16: 8e-2}[dtype]
        err = (gi.cpu().contiguous().double() - want_gi).abs().max().item()
        assert err <= tol * max(want_gi.abs().max().item(), 1.0), (dtype, err)


def test_backward_honours_grad_strides_and_empty(capi):
    g = torch.Generator().manual_seed(2)
    x = torch.randn(2, 8, 20, 20, generator=g)
    rois = torch.tensor([[0, 2.0, 2.0, 30.0, 40.0], [1, 10.0, 5.0, 70.0, 60.0], [1, -50.0, -50.0, -20.0, -20.0]])
    out, am = capi.roi_pool_forward(x.to(DEV), rois.to(DEV), 4, 6, 0.25)
    assert torch.equal(out[2].cpu(), torch.zeros(8, 4, 6)) and (am[2] == -1).all()  # empty bins: 0, nothing back
    go = torch.randn(3, 8, 6, 4, generator=g).permute(0, 1, 3, 2)  # non-contiguous
    gi = capi.roi_pool_backward(go.to(DEV), rois.to(DEV), am, 4, 6, 2, 8, 20, 20)
    want = oracle.roi_pool_backward(go, rois, am.cpu(), 4, 6, 2, 8, 20, 20)
    torch.testing.assert_close(gi.cpu(), want, rtol=1e-5, atol=1e-6)
    o0, a0 = capi.roi_pool_forward(x.to(DEV), rois[:0].to(DEV), 4, 6, 0.25)
    assert o0.shape == (0, 8, 4, 6)
    assert torch.equal(capi.roi_pool_backward(go[:0].to(DEV), rois[:0].to(DEV), a0, 4, 6, 2, 8, 20, 20).cpu(),
                       torch.zeros(2, 8, 20, 20))


@pytest.mark.parametrize("channels_last", [False, True])
def test_dispatcher_module_and_autograd(golden, channels_last):
    """torch.ops.torchvision.roi_pool through the shim, the RoIPool mirror, autograd, list-of-boxes input."""
    from cpu_vision_b200 import RoIPool

    d = golden("roi_pool.pt")
    x, rois = d["x"], d["rois"]
    c = d["cases"][0]
    xd = _mf(x.to(DEV), channels_last).requires_grad_(True)
    m = RoIPool((c["ph"], c["pw"]), c["scale"])
    assert repr(m) == f"RoIPool(output_size=({c['ph']}, {c['pw']}), spatial_scale={c['scale']})"
    y = m(xd, rois.to(DEV))
    assert torch.equal(y.detach().cpu().contiguous(), c["out"])
    y.backward(c["grad_out"].to(DEV))
    torch.testing.assert_close(xd.grad.cpu().contiguous(), c["grad_in"], rtol=1e-5, atol=1e-5)
    boxes = [rois[rois[:, 0] == i, 1:].to(DEV) for i in range(2)]
    order = torch.cat([torch.where(rois[:, 0] == i)[0] for i in range(2)])
    y2 = m(xd.detach(), boxes)
    assert torch.equal(y2.cpu().contiguous(), c["out"][order])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(x, rois)
