"""Parity of the sm_100a roi_align path against the oracle, through the C-ABI
(cpu_vision_b200.capi) and through the dispatcher (torch.ops.torchvision.*).

Tolerances (BASELINE.json north_star / SURVEY.md section 8c):
  forward fp32  : 1e-5 relative (+ 1e-6 * max|x| absolute for outputs near zero)
  forward bf16  : 1e-2 relative vs the reference's autocast semantics
                  ref_fp32(x.float(), rois.float()).to(bf16); fp16: 4e-3 absolute
                  (test/test_ops.py:130-134)
  backward fp32 : atomic reordering only -- 1e-4 relative to the fp64 truth's scale
  backward bf16 : 5e-2 of the gradient's max (every packed bf16 atomic rounds), fp16 2e-2
"""
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def close_fwd(got, want, rtol=1e-5, scale=None):
    scale = want.abs().max().item() if scale is None else scale
    torch.testing.assert_close(got.cpu().float().contiguous(), want.float(), rtol=rtol, atol=1e-6 * max(scale, 1e-30))


def close_bwd(got, want64, rtol=1e-4):
    want = want64.double()
    err = (got.cpu().double().contiguous() - want).abs().max().item()
    assert err <= rtol * max(want.abs().max().item(), 1e-30), (err, want.abs().max().item())


def rand_rois(g, k, n, w, h, lo=2.0, hi=None, integer_batch=True):
    hi = hi or min(w, h)
    cx, cy = torch.rand(k, generator=g) * w, torch.rand(k, generator=g) * h
    lg = lambda: torch.exp(torch.rand(k, generator=g) * (torch.log(torch.tensor(hi / lo))) + torch.log(torch.tensor(lo)))
    bw, bh = lg(), lg()
    x1, y1 = (cx - bw / 2).clamp(0, w - 1), (cy - bh / 2).clamp(0, h - 1)
    x2, y2 = torch.maximum((cx + bw / 2).clamp(0, w), x1 + 1), torch.maximum((cy + bh / 2).clamp(0, h), y1 + 1)
    b = torch.randint(0, n, (k,), generator=g).float().sort()[0]
    return torch.stack([b, x1, y1, x2, y2], 1)


@pytest.fixture(scope="module")
def capi():
    from cpu_vision_b200 import capi as c

    return c


# ------------------------------------------------------------------ golden fixtures (from the reference)
@pytest.mark.parametrize("channels_last", [False, True])
def test_golden_roi_align(golden, capi, channels_last):
    d = golden("roi_align.pt")
    x, rois = d["x"], d["rois"]
    N, C, H, W = x.shape
    xd = x.to(DEV)
    if channels_last:
        xd = xd.contiguous(memory_format=torch.channels_last)
    layout = capi.NHWC if channels_last else capi.NCHW
    for c in d["cases"]:
        out = capi.roi_align_forward(xd, rois.to(DEV), c["ph"], c["pw"], c["scale"], c["sr"], c["aligned"])
        close_fwd(out, c["out"])
        go = c["grad_out"].to(DEV)
        if channels_last:
            go = go.contiguous(memory_format=torch.channels_last)
        gi = capi.roi_align_backward(go, rois.to(DEV), c["scale"], c["ph"], c["pw"], N, C, H, W, c["sr"], c["aligned"],
                                     layout=layout)
        want64 = oracle.roi_align_backward(c["grad_out"].double(), rois.double(), c["scale"], c["ph"], c["pw"], N, C,
                                           H, W, c["sr"], c["aligned"])
        close_bwd(gi, want64)
        close_bwd(gi, c["grad_in"], rtol=2e-5)  # the reference's own fp32 result


def test_known_answers(golden, capi):
    k = golden("known_answers.pt")
    f = lambda x, r, ph, s, sr, al: capi.roi_align_forward(x.to(DEV), r.to(DEV), ph, ph, s, sr, al).cpu()
    close_fwd(f(k["K2_x"], k["K2_roi"], 2, 1.0, 2, False), k["K2_unaligned_sr2"])
    close_fwd(f(k["K2_x"], k["K2_roi"], 2, 1.0, -1, False), k["K2_unaligned_srm1"])
    close_fwd(f(k["K2_x"], k["K2_roi"], 2, 1.0, 2, True), k["K2_aligned_sr2"])
    close_fwd(f(k["K3_x"], k["K3_roi"], 2, 2.5, 0, True), k["K3_a"])
    close_fwd(f(k["K3_x"], k["K3_roi"], 5, 0.5, 3, True), k["K3_b"])
    close_fwd(f(k["K3_x"], k["K4_roi"], 2, 1.0, 1, True), k["K4_aligned"])
    close_fwd(f(k["K3_x"], k["K4_roi"], 2, 1.0, 1, False), k["K4_unaligned"])
    g = capi.roi_align_backward(torch.ones(1, 1, 2, 2, device=DEV), k["K2_roi"].to(DEV), 1.0, 2, 2, 1, 1, 8, 8, 2, True)
    close_bwd(g, k["K5_grad"], rtol=1e-6)
    rois = torch.cat([torch.zeros(len(k["K6_boxes"]), 1), k["K6_boxes"]], 1)
    assert torch.equal(capi.fpn_level_map(rois.to(DEV), 2, 5).cpu().long(), k["K6_levels"])


# ------------------------------------------------------------------ random parity vs the oracle
CASES = [
    # (N, C, H, W, K, PH, PW, scale, sr, aligned)
    (2, 256, 50, 68, 64, 7, 7, 0.0625, 2, True),      # box head on an FPN level, fast path (C % 128 == 0)
    (2, 256, 50, 68, 48, 14, 14, 0.0625, 2, False),   # mask head
    (1, 64, 40, 40, 32, 7, 7, 0.25, 2, True),         # C not a multiple of the warp tile
    (2, 19, 23, 31, 40, 5, 3, 0.5, 2, False),         # odd sizes
    (2, 32, 30, 30, 24, 7, 7, 0.25, -1, True),        # adaptive sampling (test_ops.py default)
    (2, 32, 30, 30, 24, 4, 4, 0.25, 3, False),        # sampling_ratio 3
    (1, 8, 10, 10, 4, 5, 5, 1.0, -1, False),          # the RoIOpTester shape
    (3, 128, 25, 34, 50, 7, 7, 0.03125, 2, False),    # coarsest FPN level, RoIs smaller than a pixel
    (2, 256, 20, 20, 16, 1, 1, 0.25, 2, True),        # 1x1 pooling
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("channels_last", [False, True])
def test_forward_backward_fp32(capi, case, channels_last):
    N, C, H, W, K, PH, PW, scale, sr, aligned = case
    g = torch.Generator().manual_seed(hash(case) % (2 ** 31))
    x = torch.randn(N, C, H, W, generator=g)
    rois = rand_rois(g, K, N, W / scale, H / scale, lo=1.0 / scale)
    want = oracle.roi_align_forward(x, rois, PH, PW, scale, sr, aligned)
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last) if channels_last else x.to(DEV)
    out = capi.roi_align_forward(xd, rois.to(DEV), PH, PW, scale, sr, aligned)
    assert out.shape == want.shape
    close_fwd(out, want)
    go = torch.randn(K, C, PH, PW, generator=g)
    want64 = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PW, N, C, H, W, sr, aligned)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last) if channels_last else go.to(DEV)
    gi = capi.roi_align_backward(god, rois.to(DEV), scale, PH, PW, N, C, H, W, sr, aligned,
                                 layout=capi.NHWC if channels_last else capi.NCHW)
    close_bwd(gi, want64)


@pytest.mark.parametrize("dtype,rtol,atol,btol", [(torch.bfloat16, 1e-2, 1e-2, 5e-2), (torch.float16, 0.0, 4e-3, 2e-2)])
@pytest.mark.parametrize("channels_last", [False, True])
@pytest.mark.parametrize("ph", [7, 14])
def test_forward_backward_16bit(capi, dtype, rtol, atol, btol, channels_last, ph):
    """Oracle = the reference's autocast semantics: fp32 compute on the widened inputs, cast back."""
    g = torch.Generator().manual_seed(11)
    N, C, H, W, K, scale = 2, 256, 50, 68, 40, 0.0625
    x = torch.randn(N, C, H, W, generator=g).to(dtype)
    rois = rand_rois(g, K, N, W / scale, H / scale, lo=16.0)
    want = oracle.roi_align_forward(x.float(), rois, ph, ph, scale, 2, True).to(dtype)
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last) if channels_last else x.to(DEV)
    out = capi.roi_align_forward(xd, rois.to(DEV), ph, ph, scale, 2, True)   # fp32 rois, 16-bit features
    assert out.dtype == dtype
    torch.testing.assert_close(out.cpu().float().contiguous(), want.float(), rtol=rtol, atol=atol)
    go = torch.randn(K, C, ph, ph, generator=g).to(dtype)
    want64 = oracle.roi_align_backward(go.double(), rois.double(), scale, ph, ph, N, C, H, W, 2, True)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last) if channels_last else go.to(DEV)
    gi = capi.roi_align_backward(god, rois.to(DEV), scale, ph, ph, N, C, H, W, 2, True,
                                 layout=capi.NHWC if channels_last else capi.NCHW)
    assert gi.dtype == dtype
    # packed 16-bit red.global.add rounds after EVERY accumulation (the reference under autocast
    # accumulates in fp32 and rounds once): n contributions cost up to ~sqrt(n) * 2^-9 (bf16) / 2^-11 (fp16).
    # The NCHW fallback kernel issues one 16-bit atomic per sample tap (16 per output element, in a
    # run-dependent order) where the channels_last kernel sums a whole column in fp32 first, so its
    # rounding error is larger and varies from run to run: twice the bound there.
    close_bwd(gi, want64, rtol=btol if channels_last else 2 * btol)


def test_fp64_matches_oracle(capi):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1, 8, 5, 5, generator=g, dtype=torch.float64)
    rois = torch.tensor([[0, 0, 0, 4, 4], [0, 0, 2, 3, 4], [0, 1.5, 0.7, 2.2, 3.9]], dtype=torch.float64)
    for sr, aligned in ((1, False), (2, True), (-1, True)):
        want = oracle.roi_align_forward(x, rois, 2, 2, 1.0, sr, aligned)
        out = capi.roi_align_forward(x.to(DEV), rois.to(DEV), 2, 2, 1.0, sr, aligned)
        torch.testing.assert_close(out.cpu(), want, rtol=1e-12, atol=1e-13)
        go = torch.randn(3, 8, 2, 2, generator=g, dtype=torch.float64)
        gi = capi.roi_align_backward(go.to(DEV), rois.to(DEV), 1.0, 2, 2, 1, 8, 5, 5, sr, aligned)
        torch.testing.assert_close(gi.cpu(), oracle.roi_align_backward(go, rois, 1.0, 2, 2, 1, 8, 5, 5, sr, aligned),
                                   rtol=1e-12, atol=1e-13)


def test_fp64_pyramid_and_single_level_levels(capi):
    """fp64 features through the FPN path: the level map runs in fp64 like LevelMapper on fp64 boxes
    (poolers.py:73-84), and a one-level pyramid returns all-zero level ids (they are saved for the backward)."""
    g = torch.Generator().manual_seed(15)
    shapes, scales = [(24, 32), (12, 16)], [0.25, 0.125]
    feats = [torch.randn(2, 4, h, w, generator=g, dtype=torch.float64) for h, w in shapes]
    rois = rand_rois(g, 40, 2, 128.0, 96.0, lo=4.0).double()
    lv = oracle.fpn_level_map(rois.float(), 2, 3)
    want = torch.zeros(40, 4, 3, 3, dtype=torch.float64)
    for l in range(2):
        sel = (lv == l).nonzero().flatten()
        want[sel] = oracle.roi_align_forward(feats[l], rois[sel], 3, 3, scales[l], 2, False)
    out, levels = capi.msroi_align_forward([f.to(DEV) for f in feats], rois.to(DEV), scales, 3, 3, 2, False, k_min=2,
                                           k_max=3)
    assert torch.equal(levels.cpu().long(), lv)
    torch.testing.assert_close(out.cpu(), want, rtol=1e-12, atol=1e-13)
    out1, lv1 = capi.msroi_align_forward([feats[0].to(DEV)], rois.to(DEV), scales[:1], 3, 3, 2, False, k_min=2, k_max=2)
    assert lv1.shape == (40,) and int(lv1.abs().sum()) == 0
    torch.testing.assert_close(out1.cpu(), oracle.roi_align_forward(feats[0], rois, 3, 3, 0.25, 2, False),
                               rtol=1e-12, atol=1e-13)


def test_edge_rois(capi):
    """Malformed, out-of-image, zero-size and sub-pixel RoIs; empty inputs."""
    g = torch.Generator().manual_seed(2)
    x = torch.randn(2, 128, 16, 20, generator=g)
    rois = torch.tensor([
        [0, 10, 12, 4, 3],        # x2 < x1, y2 < y1
        [1, -50, -50, -20, -20],  # fully outside (top-left): every sample empty
        [0, 100, 100, 150, 150],  # fully outside (bottom-right)
        [1, -3, -3, 6, 6],        # straddles the corner
        [0, 7, 7, 7, 7],          # zero size
        [1, 5.2, 5.3, 5.4, 5.6],  # much smaller than a pixel
        [0, 0, 0, 20, 16],        # the whole map
        [1, 18.5, 14.5, 25, 19],  # hugs the far border (x_low >= W-1 clamp)
    ], dtype=torch.float32)
    for layout_cl in (False, True):
        xd = x.to(DEV).contiguous(memory_format=torch.channels_last) if layout_cl else x.to(DEV)
        for sr in (2, -1):
            for aligned in (True, False):
                want = oracle.roi_align_forward(x, rois, 7, 7, 1.0, sr, aligned)
                out = capi.roi_align_forward(xd, rois.to(DEV), 7, 7, 1.0, sr, aligned)
                close_fwd(out, want, scale=x.abs().max().item())
                go = torch.randn(8, 128, 7, 7, generator=g)
                want64 = oracle.roi_align_backward(go.double(), rois.double(), 1.0, 7, 7, 2, 128, 16, 20, sr, aligned)
                god = go.to(DEV).contiguous(memory_format=torch.channels_last) if layout_cl else go.to(DEV)
                gi = capi.roi_align_backward(god, rois.to(DEV), 1.0, 7, 7, 2, 128, 16, 20, sr, aligned,
                                             layout=capi.NHWC if layout_cl else capi.NCHW)
                close_bwd(gi, want64)
    # empty RoI list / empty channel dim
    out = capi.roi_align_forward(x.to(DEV), torch.zeros(0, 5, device=DEV), 7, 7, 1.0, 2, True)
    assert out.shape == (0, 128, 7, 7)
    gi = capi.roi_align_backward(torch.zeros(0, 128, 7, 7, device=DEV), torch.zeros(0, 5, device=DEV), 1.0, 7, 7, 2,
                                 128, 16, 20, 2, True)
    assert gi.shape == (2, 128, 16, 20) and gi.abs().sum().item() == 0.0


def test_backward_honours_grad_strides(capi):
    """The reference reads grad through its strides (cuda/roi_align_kernel.cu:253-256)."""
    g = torch.Generator().manual_seed(9)
    N, C, H, W, K = 2, 128, 20, 24, 20
    rois = rand_rois(g, K, N, W * 4.0, H * 4.0, lo=8.0)
    base = torch.randn(K, 7, 7, C, generator=g)
    variants = {
        "nhwc_view": base.permute(0, 3, 1, 2),
        "expanded": torch.randn(K, C, 1, 1, generator=g).expand(K, C, 7, 7),      # e.g. grad of .mean((2,3))
        "sliced": torch.randn(K, C, 7, 14, generator=g)[:, :, :, ::2],
    }
    for name, go in variants.items():
        want64 = oracle.roi_align_backward(go.double().contiguous(), rois.double(), 0.25, 7, 7, N, C, H, W, 2, True)
        god = torch.empty_strided(go.shape, go.stride(), dtype=go.dtype, device=DEV) if name != "expanded" else None
        if god is None:
            god = go[:, :, :1, :1].contiguous().to(DEV).expand(K, C, 7, 7)
        else:
            god.copy_(go)
        assert god.stride() == go.stride(), name
        for layout in (capi.NCHW, capi.NHWC):
            gi = capi.roi_align_backward(god, rois.to(DEV), 0.25, 7, 7, N, C, H, W, 2, True, layout=layout)
            close_bwd(gi, want64)


# ------------------------------------------------------------------ MultiScaleRoIAlign
@pytest.mark.parametrize("channels_last", [False, True])
def test_msroi_align_golden(golden, capi, channels_last):
    d = golden("msroi_align.pt")
    boxes = d["boxes"]
    rois = torch.cat([torch.cat([torch.full((len(b), 1), float(i)), b], 1) for i, b in enumerate(boxes)])
    mf = torch.channels_last if channels_last else torch.contiguous_format
    feats = [f.to(DEV).contiguous(memory_format=mf) for f in d["feats"]]
    for osz in (7, 14):
        out, levels = capi.msroi_align_forward(feats, rois.to(DEV), d["scales"], osz, osz, 2, False, d["k_min"],
                                               d["k_max"])
        assert torch.equal(levels.cpu().long(), d["levels"]), "FPN levels must match exactly"
        close_fwd(out, d[f"out_{osz}"])
        go = d[f"grad_out_{osz}"].to(DEV).contiguous(memory_format=mf)
        gis = capi.msroi_align_backward(go, rois.to(DEV), levels, d["scales"], [tuple(f.shape) for f in feats], osz,
                                        osz, 2, False, layout=capi.NHWC if channels_last else capi.NCHW)
        want = oracle.msroi_align_backward(d[f"grad_out_{osz}"], rois, d["levels"], d["scales"],
                                           [tuple(f.shape) for f in feats], osz, osz, 2, False)
        for a, b, c in zip(gis, d[f"grad_in_{osz}"], want):
            close_bwd(a, b, rtol=2e-5)
            close_bwd(a, c, rtol=2e-5)


def test_msroi_module_matches_reference_module(golden):
    """cpu_vision_b200.MultiScaleRoIAlign(...)(features, boxes, image_shapes) == the reference module's output."""
    from collections import OrderedDict

    from cpu_vision_b200 import MultiScaleRoIAlign

    d = golden("msroi_align.pt")
    for cl in (False, True):
        mf = torch.channels_last if cl else torch.contiguous_format
        feats = OrderedDict((str(i), f.to(DEV).contiguous(memory_format=mf).requires_grad_(True))
                            for i, f in enumerate(d["feats"]))
        feats["pool"] = torch.zeros(2, 3, 4, 5, device=DEV)  # filtered out by featmap_names
        for osz in (7, 14):
            m = MultiScaleRoIAlign(["0", "1", "2", "3"], osz, 2)
            y = m(feats, [b.to(DEV) for b in d["boxes"]], d["image_shapes"])
            assert m.scales == d["scales"] and (m.map_levels.k_min, m.map_levels.k_max) == (d["k_min"], d["k_max"])
            close_fwd(y, d[f"out_{osz}"])
            y.backward(d[f"grad_out_{osz}"].to(DEV))
            for i in range(4):
                close_bwd(feats[str(i)].grad, d[f"grad_in_{osz}"][i], rtol=2e-5)
                feats[str(i)].grad = None
        assert torch.equal(m.map_levels([b.to(DEV) for b in d["boxes"]]).cpu(), d["levels"])


# ------------------------------------------------------------------ dispatcher (the drop-in boundary)
def test_dispatcher_ops_and_autograd():
    import cpu_vision_b200 as m

    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 64, 24, 24, generator=g)
    rois = rand_rois(g, 30, 2, 96.0, 96.0, lo=6.0)
    want = oracle.roi_align_forward(x, rois, 7, 7, 0.25, 2, True)
    xd = x.to(DEV).requires_grad_(True)
    y = torch.ops.torchvision.roi_align(xd, rois.to(DEV), 0.25, 7, 7, 2, True)
    close_fwd(y, want)
    go = torch.randn(30, 64, 7, 7, generator=g)
    y.backward(go.to(DEV))
    close_bwd(xd.grad, oracle.roi_align_backward(go.double(), rois.double(), 0.25, 7, 7, 2, 64, 24, 24, 2, True))
    # list-of-boxes front-end, channels_last in -> channels_last out
    xcl = x.to(DEV).contiguous(memory_format=torch.channels_last)
    boxes = [rois[rois[:, 0] == i, 1:].to(DEV) for i in range(2)]
    y2 = m.roi_align(xcl, boxes, 7, 0.25, 2, True)
    assert y2.is_contiguous(memory_format=torch.channels_last)
    close_fwd(y2, want)
    assert torch.equal(m.RoIAlign(7, 0.25, 2, True)(xcl, boxes), y2)
    # permuted (non-contiguous) input, like test_ops.py:142-143
    xp = x.permute(0, 1, 3, 2).contiguous().permute(0, 1, 3, 2).to(DEV)
    assert not xp.is_contiguous()
    close_fwd(torch.ops.torchvision.roi_align(xp, rois.to(DEV), 0.25, 7, 7, 2, True), want)
    # error behaviour of cuda/roi_align_kernel.cu:342-350
    with pytest.raises(RuntimeError):
        torch.ops.torchvision.roi_align(xd, rois[:, :4].to(DEV), 0.25, 7, 7, 2, True)
    with pytest.raises(RuntimeError):
        torch.ops.torchvision.roi_align(xd, rois.to(DEV).double(), 0.25, 7, 7, 2, True)
    with pytest.raises(RuntimeError):
        torch.ops.torchvision.roi_align(xd, rois, 0.25, 7, 7, 2, True)  # rois on the CPU


def test_gradcheck_fp64():
    """test_ops.py:189-213: gradcheck on rand(1, 8, 5, 5), pool 2, sampling_ratio 1."""
    import cpu_vision_b200  # noqa: F401

    torch.manual_seed(0)
    x = torch.rand(1, 8, 5, 5, dtype=torch.float64, device=DEV, requires_grad=True)
    rois = torch.tensor([[0, 0, 0, 4, 4], [0, 0, 2, 3, 4], [0, 2, 2, 4, 4]], dtype=torch.float64, device=DEV)
    f = lambda z: torch.ops.torchvision.roi_align(z, rois, 1.0, 2, 2, 1, False)
    assert torch.autograd.gradcheck(f, (x,), nondet_tol=1e-5)
    xp = x.detach().permute(0, 1, 3, 2).contiguous().permute(0, 1, 3, 2).requires_grad_(True)
    assert torch.autograd.gradcheck(f, (xp,), nondet_tol=1e-5)


# ------------------------------------------------------------------ full-size properties (BASELINE configs)
def test_cfg1_full_size_adjoint_and_sample(capi):
    """cfg1: 2x256x200x272, 1000 RoIs/img, 7x7, sr=2, aligned.  The oracle checks a
    slice; the whole result is checked through the adjoint identity
    <roi_align(x), g> == <x, roi_align_backward(g)> which holds for any size."""
    g = torch.Generator().manual_seed(1234 + 1)
    N, C, H, W, K = 2, 256, 200, 272, 2000
    x = torch.randn(N, C, H, W, generator=g)
    rois = rand_rois(g, K, N, 1088.0, 800.0, lo=16.0, hi=512.0)
    for cl in (True, False):
        mf = torch.channels_last if cl else torch.contiguous_format
        xd = x.to(DEV).contiguous(memory_format=mf)
        out = capi.roi_align_forward(xd, rois.to(DEV), 7, 7, 0.25, 2, True)
        go = torch.randn(K, C, 7, 7, generator=g).to(DEV).contiguous(memory_format=mf)
        gi = capi.roi_align_backward(go, rois.to(DEV), 0.25, 7, 7, N, C, H, W, 2, True,
                                     layout=capi.NHWC if cl else capi.NCHW)
        lhs = (out.double() * go.double()).sum().item()
        rhs = (xd.double() * gi.double()).sum().item()
        assert abs(lhs - rhs) <= 1e-6 * max(abs(lhs), (out.double().abs() * go.double().abs()).sum().item()), (lhs, rhs)
        sel = torch.arange(0, K, 97)
        want = oracle.roi_align_forward(x[:, :16], rois[sel], 7, 7, 0.25, 2, True)
        close_fwd(out[sel][:, :16], want)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_cfg3_full_size_levels_and_adjoint(capi, dtype):
    """cfg3: 4 FPN levels of an 800x1088 image, batch 16, 512 RoIs/img, 14x14, fwd+bwd."""
    g = torch.Generator().manual_seed(1234 + 3)
    B, C, PH = 16, 256, 14
    shapes = [(200, 272), (100, 136), (50, 68), (25, 34)]
    scales = [0.25, 0.125, 0.0625, 0.03125]
    feats = [torch.randn(B, C, h, w, generator=g).to(dtype).to(DEV).contiguous(memory_format=torch.channels_last)
             for h, w in shapes]
    rois = rand_rois(g, B * 512, B, 1088.0, 800.0, lo=16.0, hi=800.0)
    out, levels = capi.msroi_align_forward(feats, rois.to(DEV), scales, PH, PH, 2, False)
    want_lv = oracle.fpn_level_map(rois, 2, 5)
    assert torch.equal(levels.cpu().long(), want_lv)
    assert want_lv.bincount(minlength=4).min().item() > 0
    go = torch.randn(B * 512, C, PH, PH, generator=g).to(dtype).to(DEV).contiguous(memory_format=torch.channels_last)
    gis = capi.msroi_align_backward(go, rois.to(DEV), levels, scales, [(B, C, h, w) for h, w in shapes], PH, PH, 2,
                                    False, layout=capi.NHWC)
    lhs = (out.double() * go.double()).sum().item()
    rhs = sum((f.double() * gi.double()).sum().item() for f, gi in zip(feats, gis))
    mag = (out.double().abs() * go.double().abs()).sum().item()
    assert abs(lhs - rhs) <= (1e-6 if dtype == torch.float32 else 5e-3) * mag, (lhs, rhs, mag)
    # a sample of RoIs against the oracle (autocast semantics for bf16)
    sel = torch.arange(0, B * 512, 509)
    want = oracle.msroi_align_forward([f.float().cpu()[:, :8].contiguous() for f in feats], rois[sel], want_lv[sel],
                                      scales, PH, PH, 2, False)
    got = out[sel.to(DEV)][:, :8].float().cpu().contiguous()
    if dtype == torch.float32:
        close_fwd(got, want)
    else:
        torch.testing.assert_close(got, want.to(dtype).float(), rtol=1e-2, atol=1e-2)


# ------------------------------------------------------------------ backward variants (C-ABI flags)
def _bwd_case(seed, N=3, C=256, H=40, W=56, K=200, PH=7, scale=0.25, lo=8.0, sort=True):
    g = torch.Generator().manual_seed(seed)
    rois = rand_rois(g, K, N, W / scale, H / scale, lo=lo)
    if not sort:
        rois = rois[torch.randperm(K, generator=g)]
    go = torch.randn(K, C, PH, PH, generator=g)
    want64 = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PH, N, C, H, W, 2, False)
    return rois, go, want64, (N, C, H, W, PH, scale)


@pytest.mark.parametrize("ph,scale,lo", [(7, 0.25, 8.0), (14, 0.25, 16.0), (7, 0.0625, 4.0)])
def test_backward_deterministic_kernel(capi, ph, scale, lo):
    """B200DET_BWD_DETERMINISTIC: strip accumulation in shared memory, no atomics -- bitwise reproducible and equal to the
    fp64 oracle within fp32 rounding (the role of ops/roi_align.py:251-255 in the reference)."""
    rois, go, want64, (N, C, H, W, PH, sc) = _bwd_case(11 + ph, PH=ph, scale=scale, lo=lo)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    runs = [capi.roi_align_backward(god, rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False, layout=capi.NHWC,
                                    deterministic=True) for _ in range(3)]
    close_bwd(runs[0], want64)
    assert torch.equal(runs[0], runs[1]) and torch.equal(runs[0], runs[2])
    # bf16 I/O, fp32 accumulation in shared memory: one rounding at the end, much tighter than packed atomics
    gb = capi.roi_align_backward(god.to(torch.bfloat16), rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False,
                                 layout=capi.NHWC, deterministic=True)
    want_b = oracle.roi_align_backward(go.to(torch.bfloat16).double(), rois.double(), sc, PH, PH, N, C, H, W, 2, False)
    close_bwd(gb.float(), want_b, rtol=1e-2)
    # not available for NCHW: the call must say so instead of silently using atomics
    with pytest.raises(RuntimeError):
        capi.roi_align_backward(go.to(DEV), rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False, layout=capi.NCHW,
                                deterministic=True)


@pytest.mark.parametrize("deterministic", [False, True])
def test_backward_rois_not_grouped_by_image(capi, deterministic):
    """Tensor[K,5] rois may come in any order (roi_align_kernel.cu:90 just reads the batch index)."""
    rois, go, want64, (N, C, H, W, PH, sc) = _bwd_case(5, sort=False)
    assert not bool((rois[1:, 0] >= rois[:-1, 0]).all())
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    gi = capi.roi_align_backward(god, rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False, layout=capi.NHWC,
                                 deterministic=deterministic)
    close_bwd(gi, want64)


@pytest.mark.parametrize("deterministic", [False, True])
def test_backward_accumulates_without_zero_init(capi, deterministic):
    rois, go, want64, (N, C, H, W, PH, sc) = _bwd_case(6)
    g = torch.Generator().manual_seed(60)
    base = torch.randn(N, C, H, W, generator=g)
    out = base.to(DEV).contiguous(memory_format=torch.channels_last)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    gi = capi.roi_align_backward(god, rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False, layout=capi.NHWC, out=out,
                                 zero_init=False, deterministic=deterministic)
    assert gi.data_ptr() == out.data_ptr()
    close_bwd(gi, want64 + base.double())


def test_deterministic_mode_through_the_dispatcher():
    """torch.use_deterministic_algorithms(True): channels_last gradients take the tile kernel (no error, bitwise
    equal runs); NCHW ones raise like the reference's alertNotDeterministic (roi_align_kernel.cu:440)."""
    import cpu_vision_b200  # noqa: F401

    rois, go, want64, (N, C, H, W, PH, sc) = _bwd_case(7)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    torch.use_deterministic_algorithms(True)
    try:
        a = torch.ops.torchvision._roi_align_backward(god, rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False)
        b = torch.ops.torchvision._roi_align_backward(god, rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False)
        assert torch.equal(a, b)
        close_bwd(a, want64)
        # an NCHW gradient that qualifies for the fast kernels is re-laid out and takes the tile kernel too
        c = torch.ops.torchvision._roi_align_backward(go.to(DEV), rois.to(DEV), sc, PH, PH, N, C, H, W, 2, False)
        assert c.is_contiguous() and torch.equal(c, a.contiguous())
        with pytest.raises(RuntimeError):  # C = 3: generic atomics kernel, alert like the reference
            torch.ops.torchvision._roi_align_backward(go[:, :3].contiguous().to(DEV), rois.to(DEV), sc, PH, PH, N, 3,
                                                      H, W, 2, False)
    finally:
        torch.use_deterministic_algorithms(False)


@pytest.mark.parametrize("chunks", [None, "1", "3"])
@pytest.mark.parametrize("channels_last", [False, True])
def test_msroi_align_host_entry_points_pipelined(golden, capi, channels_last, chunks, monkeypatch):
    """b200det_msroi_align_{forward,backward}_host: CPU tensors in/out, pipelined over image chunks
    (H2D / kernels / D2H overlap); results must equal the reference fixtures and the device entry points.
    B200DET_HOST_CHUNKS (read per call) sets the number of chunks: default (one per image here), 1 (no
    pipelining) and 3 (uneven chunks) must all give the same answer."""
    if chunks is None:
        monkeypatch.delenv("B200DET_HOST_CHUNKS", raising=False)
    else:
        monkeypatch.setenv("B200DET_HOST_CHUNKS", chunks)
    d = golden("msroi_align.pt")
    boxes = d["boxes"]
    rois = torch.cat([torch.cat([torch.full((len(b), 1), float(i)), b], 1) for i, b in enumerate(boxes)])
    mf = torch.channels_last if channels_last else torch.contiguous_format
    layout = capi.NHWC if channels_last else capi.NCHW
    feats = [f.contiguous(memory_format=mf).pin_memory() for f in d["feats"]]
    shapes = [tuple(f.shape) for f in feats]
    assert shapes[0][0] >= 2, "fixture must have several images so that the call really is chunked"
    out, levels = capi.msroi_align_forward_host(feats, rois, d["scales"], 7, 7, 2, False, d["k_min"], d["k_max"])
    assert torch.equal(levels.long(), d["levels"])
    close_fwd(out, d["out_7"])
    go = d["grad_out_7"].contiguous(memory_format=mf).pin_memory()
    gis = capi.msroi_align_backward_host(go, rois, levels, d["scales"], shapes, 7, 7, 2, False, layout=layout)
    for a, b in zip(gis, d["grad_in_7"]):
        close_bwd(a, b, rtol=2e-5)
    # RoIs in arbitrary image order: the call falls back to one chunk and must still be right
    perm = torch.randperm(rois.shape[0], generator=torch.Generator().manual_seed(3))
    out_p, lv_p = capi.msroi_align_forward_host(feats, rois[perm], d["scales"], 7, 7, 2, False, d["k_min"], d["k_max"])
    close_fwd(out_p, d["out_7"][perm])
    gis_p = capi.msroi_align_backward_host(go[perm].contiguous(memory_format=mf), rois[perm], lv_p, d["scales"], shapes,
                                           7, 7, 2, False, layout=layout)
    for a, b in zip(gis_p, d["grad_in_7"]):
        close_bwd(a, b, rtol=2e-5)
    capi.host_release()


# ---- NCHW callers (the reference's default layout): relayout -> channels_last kernels -> relayout ---------
@pytest.mark.parametrize("pipe", ["1", "0"])   # B200DET_RELAYOUT_PIPE: pipelined cp.async kernel / load-barrier-store kernels
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16, torch.float64])
@pytest.mark.parametrize("shape", [(1, 1, 1, 1), (2, 3, 5, 7), (3, 256, 40, 56), (2, 130, 33, 65), (1, 64, 1, 200),
                                   (2, 200, 24, 50), (70, 256, 14, 14), (40, 128, 7, 8), (300, 256, 7, 7)])
def test_relayout_both_ways(capi, dtype, shape, pipe, monkeypatch):
    monkeypatch.setenv("B200DET_RELAYOUT_PIPE", pipe)
    g = torch.Generator().manual_seed(sum(shape))
    x = torch.randn(shape, generator=g).to(dtype).to(DEV)
    cl = capi.relayout(x, True)
    assert cl.is_contiguous(memory_format=torch.channels_last) and torch.equal(cl, x)
    back = capi.relayout(cl, False)
    assert back.is_contiguous() and torch.equal(back, x)


def test_nchw_caller_takes_the_fast_kernels_through_the_dispatcher():
    """Same tensors, NCHW-contiguous like the reference's models produce them: results are NCHW-contiguous,
    equal to the channels_last call (same kernels) and to the oracle; autograd included."""
    import cpu_vision_b200  # noqa: F401

    g = torch.Generator().manual_seed(21)
    N, C, H, W, K, PH = 2, 256, 50, 68, 300, 7
    x = torch.randn(N, C, H, W, generator=g)
    rois = rand_rois(g, K, N, W * 16.0, H * 16.0, lo=16.0)
    want = oracle.roi_align_forward(x, rois, PH, PH, 0.0625, 2, True)
    xd = x.to(DEV).requires_grad_(True)
    y = torch.ops.torchvision.roi_align(xd, rois.to(DEV), 0.0625, PH, PH, 2, True)
    assert y.is_contiguous()
    close_fwd(y.detach(), want)
    ycl = torch.ops.torchvision.roi_align(x.to(DEV).contiguous(memory_format=torch.channels_last), rois.to(DEV), 0.0625,
                                          PH, PH, 2, True)
    assert torch.equal(y.detach(), ycl.contiguous())
    go = torch.randn(K, C, PH, PH, generator=g)
    y.backward(go.to(DEV))
    assert xd.grad.is_contiguous()
    close_bwd(xd.grad, oracle.roi_align_backward(go.double(), rois.double(), 0.0625, PH, PH, N, C, H, W, 2, True))
    # a handful of RoIs on a big map: the re-layout would cost more than it saves, generic kernel, same answer
    few = rois[:2].to(DEV)
    close_fwd(torch.ops.torchvision.roi_align(x.to(DEV), few, 0.0625, PH, PH, 2, True),
              oracle.roi_align_forward(x, rois[:2], PH, PH, 0.0625, 2, True))


def test_nchw_relayout_switch(monkeypatch):
    """B200DET_NCHW_RELAYOUT=0 keeps an NCHW caller on the generic strided kernels: same values (forward within
    the forward bar, backward within the reordering bar) as the re-layout path."""
    import cpu_vision_b200  # noqa: F401

    g = torch.Generator().manual_seed(22)
    N, C, H, W, K, PH = 2, 128, 40, 56, 200, 7
    x = torch.randn(N, C, H, W, generator=g).to(DEV)
    rois = rand_rois(g, K, N, W * 8.0, H * 8.0, lo=8.0).to(DEV)
    go = torch.randn(K, C, PH, PH, generator=g).to(DEV)
    res = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("B200DET_NCHW_RELAYOUT", flag)
        xd = x.clone().requires_grad_(True)
        y = torch.ops.torchvision.roi_align(xd, rois, 0.125, PH, PH, 2, False)
        y.backward(go)
        res[flag] = (y.detach(), xd.grad)
    close_fwd(res["0"][0], res["1"][0].cpu())
    close_bwd(res["0"][1], res["1"][1].cpu().double())


def test_nchw_msroi_module_through_relayout(golden):
    """MultiScaleRoIAlign mirror on NCHW features with C = 256: b200det::msroi_align re-lays out every level."""
    from cpu_vision_b200.poolers import MultiScaleRoIAlign
    from collections import OrderedDict

    g = torch.Generator().manual_seed(5)
    img = (256, 320)
    feats = OrderedDict((str(i), torch.randn(2, 256, img[0] // (4 << i), img[1] // (4 << i), generator=g))
                        for i in range(4))
    boxes = []
    for _ in range(2):
        r = rand_rois(g, 150, 1, float(img[1]), float(img[0]), lo=8.0, hi=400.0)[:, 1:]
        boxes.append(r)
    m = MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2)
    fd = OrderedDict((k, v.to(DEV).requires_grad_(True)) for k, v in feats.items())
    y = m(fd, [b.to(DEV) for b in boxes], [img, img])
    assert y.is_contiguous()
    rois = torch.cat([torch.cat([torch.full((b.shape[0], 1), float(i)), b], 1) for i, b in enumerate(boxes)])
    levels = oracle.fpn_level_map(rois, 2, 5)
    scales = [1.0 / (4 << i) for i in range(4)]
    want = oracle.msroi_align_forward(list(feats.values()), rois, levels, scales, 7, 7, 2, False)
    close_fwd(y.detach(), want)
    go = torch.randn(y.shape, generator=g)
    y.backward(go.to(DEV))
    wants = oracle.msroi_align_backward(go, rois, levels, scales, [tuple(f.shape) for f in feats.values()], 7, 7, 2, False)
    for f, w_ in zip(fd.values(), wants):
        assert f.grad.is_contiguous()
        close_bwd(f.grad, w_)


def test_box_head_layout_fusion(capi):
    """Row f2: channels_last pooled features feed fc6 as a view once fc6's columns are permuted."""
    from cpu_vision_b200 import box_head

    g = torch.Generator().manual_seed(9)
    N, C, H, W, K, PH = 2, 256, 25, 34, 64, 7
    x = torch.randn(N, C, H, W, generator=g).to(DEV).contiguous(memory_format=torch.channels_last)
    rois = rand_rois(g, K, N, W * 32.0, H * 32.0, lo=32.0).to(DEV)
    y = capi.roi_align_forward(x, rois, PH, PH, 1 / 32, 2, False)
    flat = box_head.flatten_pooled(y)
    assert flat.data_ptr() == y.data_ptr() and flat.shape == (K, C * PH * PH)  # a view, not a copy
    w6 = torch.randn(128, C * PH * PH, generator=g).to(DEV) / 100
    want = torch.nn.functional.linear(y.contiguous().flatten(1), w6)  # the reference's order
    got = torch.nn.functional.linear(flat, box_head.fc6_weight_for_channels_last(w6, C, PH, PH))
    torch.testing.assert_close(got, want, rtol=1e-3, atol=1e-3)
    assert box_head.flatten_pooled(y.contiguous()).shape == flat.shape
