"""The oracle (oracle/detection_oracle.c) pinned against the reference:
known-answer vectors, golden fixtures generated from the reference itself
(tests/golden/make_golden.py), and -- when oracle/_ref/libtvref.so is present --
the reference's own compiled CPU kernels, bit for bit."""
import pytest
import torch

import oracle


def test_known_answers_survey_values(golden):
    """Literal values listed in SURVEY.md section 8(a), K1-K6."""
    k = golden("known_answers.pt")
    assert oracle.nms(k["K1_boxes"], k["K1_scores"], 0.2).tolist() == [1]
    o = lambda *a: oracle.roi_align_forward(*a).flatten()
    x8, r8 = k["K2_x"], k["K2_roi"]
    assert o(x8, r8, 2, 2, 1.0, 2, False).tolist() == [7.75, 9.25, 19.75, 21.25]
    assert o(x8, r8, 2, 2, 1.0, -1, False).tolist() == [7.75, 9.25, 19.75, 21.25]
    assert o(x8, r8, 2, 2, 1.0, 2, True).tolist() == [3.8125, 5.25, 15.3125, 16.75]
    x10 = k["K3_x"]
    torch.testing.assert_close(o(x10, k["K3_roi"], 2, 2, 2.5, 0, True),
                               torch.tensor([25.1875, 25.027779, 65.1875, 58.361111]), rtol=1e-6, atol=0)
    torch.testing.assert_close(o(x10, k["K3_roi"], 5, 5, 0.5, 3, True)[:5],
                               torch.tensor([0.0, 0.24500002, 0.67500007, 1.1050000, 1.5350000]), rtol=1e-6, atol=0)
    torch.testing.assert_close(o(x10, k["K4_roi"], 2, 2, 1.0, 1, True),
                               torch.tensor([2.3750002, 2.1250002, 8.375, 8.1250010]), rtol=1e-6, atol=0)
    torch.testing.assert_close(o(x10, k["K4_roi"], 2, 2, 1.0, 1, False),
                               torch.tensor([8.25, 8.75, 14.250001, 14.750001]), rtol=1e-6, atol=0)
    g = oracle.roi_align_backward(torch.ones(1, 1, 2, 2), r8, 1.0, 2, 2, 1, 1, 8, 8, 2, True)
    assert g.sum().item() == 4.0
    assert g[0, 0, :2, :2].flatten().tolist() == [0.47265625, 0.4296875, 0.4296875, 0.390625]
    sides = [10, 111, 112, 113, 223, 224, 225, 447, 448, 800, 2000]
    want = [0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 3]
    rois = torch.tensor([[0.0, 0, 0, s, s] for s in sides])
    assert oracle.fpn_level_map(rois, 2, 5).tolist() == want


def test_known_answers_reference_outputs(golden):
    """Same fixtures, compared with what the reference build produced (bit exact)."""
    k = golden("known_answers.pt")
    assert torch.equal(oracle.nms(k["K1_boxes"], k["K1_scores"], 0.2), k["K1_keep"])
    f = oracle.roi_align_forward
    assert torch.equal(f(k["K2_x"], k["K2_roi"], 2, 2, 1.0, 2, False), k["K2_unaligned_sr2"])
    assert torch.equal(f(k["K2_x"], k["K2_roi"], 2, 2, 1.0, -1, False), k["K2_unaligned_srm1"])
    assert torch.equal(f(k["K2_x"], k["K2_roi"], 2, 2, 1.0, 2, True), k["K2_aligned_sr2"])
    assert torch.equal(f(k["K3_x"], k["K3_roi"], 2, 2, 2.5, 0, True), k["K3_a"])
    assert torch.equal(f(k["K3_x"], k["K3_roi"], 5, 5, 0.5, 3, True), k["K3_b"])
    assert torch.equal(f(k["K3_x"], k["K4_roi"], 2, 2, 1.0, 1, True), k["K4_aligned"])
    assert torch.equal(f(k["K3_x"], k["K4_roi"], 2, 2, 1.0, 1, False), k["K4_unaligned"])
    g = oracle.roi_align_backward(torch.ones(1, 1, 2, 2), k["K2_roi"], 1.0, 2, 2, 1, 1, 8, 8, 2, True)
    assert torch.equal(g, k["K5_grad"])
    rois = torch.cat([torch.zeros(len(k["K6_boxes"]), 1), k["K6_boxes"]], 1)
    assert torch.equal(oracle.fpn_level_map(rois, 2, 5), k["K6_levels"])


def test_roi_align_golden(golden):
    d = golden("roi_align.pt")
    x, rois = d["x"], d["rois"]
    N, C, H, W = x.shape
    for c in d["cases"]:
        out = oracle.roi_align_forward(x, rois, c["ph"], c["pw"], c["scale"], c["sr"], c["aligned"])
        assert torch.equal(out, c["out"]), (c["ph"], c["sr"], c["aligned"])
        gi = oracle.roi_align_backward(c["grad_out"], rois, c["scale"], c["ph"], c["pw"], N, C, H, W, c["sr"],
                                       c["aligned"])
        assert torch.equal(gi, c["grad_in"]), (c["ph"], c["sr"], c["aligned"])


def test_nms_golden(golden):
    d = golden("nms.pt")
    for thr in (0.2, 0.5, 0.7):
        # thr_is_float=False is the CPU kernel's double comparison (what made the fixture)
        assert torch.equal(oracle.nms(d["boxes"], d["scores"], thr, thr_is_float=False), d[f"keep_{thr}"])
        assert torch.equal(oracle.nms(d["boxes"], d["scores"], thr, thr_is_float=True), d[f"keep_{thr}"])
    assert torch.equal(oracle.nms(d["boxes"], d["scores_ties"], 0.5), d["keep_ties"])
    assert torch.equal(oracle.batched_nms(d["boxes"], d["scores"], d["idxs"], 0.5, "trick"), d["bn_trick"])
    assert torch.equal(oracle.batched_nms(d["boxes"], d["scores"], d["idxs"], 0.5, "vanilla"), d["bn_vanilla"])
    assert torch.equal(oracle.batched_nms(d["boxes2"], d["scores2"], d["idxs2"], 0.5, "trick"), d["bn2_trick"])
    assert torch.equal(oracle.batched_nms(d["boxes2"], d["scores2"], d["idxs2"], 0.5, "vanilla"), d["bn2_vanilla"])


def test_msroi_align_golden(golden):
    d = golden("msroi_align.pt")
    assert d["K7_scales"] == [0.25, 0.125, 0.0625, 0.03125] and (d["K7_kmin"], d["K7_kmax"]) == (2, 5)
    boxes = d["boxes"]
    rois = torch.cat([torch.cat([torch.full((len(b), 1), float(i)), b], 1) for i, b in enumerate(boxes)])
    levels = oracle.fpn_level_map(rois, d["k_min"], d["k_max"])
    assert torch.equal(levels, d["levels"])
    assert len(set(levels.tolist())) == 4, "fixture should populate every level"
    for osz in (7, 14):
        out = oracle.msroi_align_forward(d["feats"], rois, levels, d["scales"], osz, osz, 2, False)
        assert torch.equal(out, d[f"out_{osz}"])
        gis = oracle.msroi_align_backward(d[f"grad_out_{osz}"], rois, levels, d["scales"],
                                          [tuple(f.shape) for f in d["feats"]], osz, osz, 2, False)
        for a, b in zip(gis, d[f"grad_in_{osz}"]):
            assert torch.equal(a, b)


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/libtvref.so not built")
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_restatement_matches_compiled_reference(dtype):
    """Random inputs, every sampling mode: restatement == reference kernels, bit for bit."""
    g = torch.Generator().manual_seed(7)
    x = torch.randn(3, 5, 17, 23, generator=g, dtype=dtype)
    K = 40
    rois = torch.rand(K, 5, generator=g, dtype=dtype) * 30 - 3
    rois[:, 0] = torch.randint(0, 3, (K,), generator=g).to(dtype)
    rois[:, 3:] = rois[:, 1:3] + torch.rand(K, 2, generator=g, dtype=dtype) * 25 - 2  # some malformed
    for sr in (1, 2, 4, 0, -1):
        for aligned in (False, True):
            for scale in (1.0, 0.5, 0.3):
                a = oracle.roi_align_forward(x, rois, 4, 6, scale, sr, aligned, impl="oracle")
                b = oracle.roi_align_forward(x, rois, 4, 6, scale, sr, aligned, impl="ref")
                assert torch.equal(a, b), (sr, aligned, scale)
                go = torch.randn(K, 5, 4, 6, generator=g, dtype=dtype)
                go_nc = go.permute(0, 1, 3, 2).contiguous().permute(0, 1, 3, 2)  # non-contiguous grad
                ga = oracle.roi_align_backward(go_nc, rois, scale, 4, 6, 3, 5, 17, 23, sr, aligned, impl="oracle")
                gb = oracle.roi_align_backward(go_nc, rois, scale, 4, 6, 3, 5, 17, 23, sr, aligned, impl="ref")
                assert torch.equal(ga, gb), (sr, aligned, scale)
    for seed in range(5):
        g.manual_seed(seed)
        b = torch.rand(700, 4, generator=g, dtype=dtype) * 100
        b[:, 2:] += b[:, :2]
        s = torch.rand(700, generator=g, dtype=dtype)
        for thr in (0.2, 0.5, 0.8):
            assert torch.equal(oracle.nms(b, s, thr, thr_is_float=False, impl="oracle"),
                               oracle.nms(b, s, thr, impl="ref"))


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/libtvref.so not built")
def test_reference_threaded_wrappers_equal_single_thread():
    g = torch.Generator().manual_seed(3)
    x = torch.randn(4, 6, 20, 20, generator=g)
    rois = torch.rand(64, 5, generator=g) * 15
    rois[:, 0] = torch.arange(64) // 16
    rois[:, 3:] += rois[:, 1:3]
    a = oracle.roi_align_forward(x, rois, 7, 7, 1.0, 2, True, impl="ref", threads=1)
    b = oracle.roi_align_forward(x, rois, 7, 7, 1.0, 2, True, impl="ref", threads=4)
    assert torch.equal(a, b)
    go = torch.randn(64, 6, 7, 7, generator=g)
    ga = oracle.roi_align_backward(go, rois, 1.0, 7, 7, 4, 6, 20, 20, 2, True, impl="ref", threads=1)
    gb = oracle.roi_align_backward(go, rois, 1.0, 7, 7, 4, 6, 20, 20, 2, True, impl="ref", threads=4)
    assert torch.equal(ga, gb)


# ---- roi_pool -------------------------------------------------------
def test_roi_pool_golden(golden):
    """Restatement vs outputs of the reference's own ops.roi_pool + autograd (tests/golden/make_golden_roi_pool.py)."""
    d = golden("roi_pool.pt")
    x, rois = d["x"], d["rois"]
    for c in d["cases"]:
        out, am = oracle.roi_pool_forward(x, rois, c["ph"], c["pw"], c["scale"])
        assert torch.equal(out, c["out"]), (c["ph"], c["scale"])
        gi = oracle.roi_pool_backward(c["grad_out"], rois, am, c["ph"], c["pw"], *x.shape)
        assert torch.equal(gi, c["grad_in"])


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/libtvref.so not built")
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_roi_pool_restatement_matches_compiled_reference(dtype):
    g = torch.Generator().manual_seed(11)
    x = torch.randn(3, 5, 17, 23, generator=g, dtype=dtype)
    K = 50
    rois = torch.rand(K, 5, generator=g, dtype=dtype) * 30 - 3
    rois[:, 0] = torch.randint(0, 3, (K,), generator=g).to(dtype)
    rois[:, 3:] = rois[:, 1:3] + torch.rand(K, 2, generator=g, dtype=dtype) * 25 - 2
    for scale in (1.0, 0.5, 0.3):
        for ph, pw in ((4, 6), (1, 1), (7, 7)):
            a, aa = oracle.roi_pool_forward(x, rois, ph, pw, scale, impl="oracle")
            b, ab = oracle.roi_pool_forward(x, rois, ph, pw, scale, impl="ref")
            assert torch.equal(a, b) and torch.equal(aa, ab), (scale, ph, pw)
            go = torch.randn(K, 5, ph, pw, generator=g, dtype=dtype)
            go_nc = go.permute(0, 1, 3, 2).contiguous().permute(0, 1, 3, 2)
            assert torch.equal(oracle.roi_pool_backward(go_nc, rois, aa, ph, pw, 3, 5, 17, 23, impl="oracle"),
                               oracle.roi_pool_backward(go_nc, rois, ab, ph, pw, 3, 5, 17, 23, impl="ref"))


def test_ps_roi_pool_golden(golden):
    d = golden("ps_roi_pool.pt")
    for c in d["cases"]:
        out, cm = oracle.ps_roi_pool_forward(c["x"], d["rois"], c["ph"], c["pw"], c["scale"])
        assert torch.equal(out, c["out"]), (c["ph"], c["scale"])
        gi = oracle.ps_roi_pool_backward(c["grad_out"], d["rois"], cm, c["scale"], c["ph"], c["pw"], *c["x"].shape)
        assert torch.equal(gi, c["grad_in"])


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/libtvref.so not built")
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_ps_roi_pool_restatement_matches_compiled_reference(dtype):
    g = torch.Generator().manual_seed(13)
    K = 50
    rois = torch.rand(K, 5, generator=g, dtype=dtype) * 30 - 3
    rois[:, 0] = torch.randint(0, 3, (K,), generator=g).to(dtype)
    rois[:, 3:] = rois[:, 1:3] + torch.rand(K, 2, generator=g, dtype=dtype) * 25 - 2
    for ph, pw, cm_ in ((2, 3, 2), (1, 1, 4), (3, 3, 1)):
        x = torch.randn(3, cm_ * ph * pw, 17, 23, generator=g, dtype=dtype)
        for scale in (1.0, 0.5, 0.3):
            a, ma = oracle.ps_roi_pool_forward(x, rois, ph, pw, scale, impl="oracle")
            b, mb = oracle.ps_roi_pool_forward(x, rois, ph, pw, scale, impl="ref")
            assert torch.equal(a, b) and torch.equal(ma, mb), (scale, ph, pw)
            go = torch.randn(a.shape, generator=g, dtype=dtype)
            assert torch.equal(oracle.ps_roi_pool_backward(go, rois, ma, scale, ph, pw, *x.shape, impl="oracle"),
                               oracle.ps_roi_pool_backward(go, rois, mb, scale, ph, pw, *x.shape, impl="ref"))


def test_ps_roi_align_golden(golden):
    d = golden("ps_roi_align.pt")
    for c in d["cases"]:
        out, cm = oracle.ps_roi_align_forward(c["x"], d["rois"], c["ph"], c["pw"], c["scale"], c["sr"])
        # adaptive grids of malformed RoIs are empty: 0 / 0 = NaN in the reference and here
        assert torch.equal(out.isnan(), c["out"].isnan()), (c["ph"], c["scale"], c["sr"])
        assert torch.equal(out.nan_to_num(), c["out"].nan_to_num()), (c["ph"], c["scale"], c["sr"])
        gi = oracle.ps_roi_align_backward(c["grad_out"], d["rois"], cm, c["scale"], c["ph"], c["pw"], c["sr"],
                                          *c["x"].shape)
        assert torch.equal(gi.nan_to_num(), c["grad_in"].nan_to_num())


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/libtvref.so not built")
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_ps_roi_align_restatement_matches_compiled_reference(dtype):
    g = torch.Generator().manual_seed(17)
    K = 50
    rois = torch.rand(K, 5, generator=g, dtype=dtype) * 30 - 3
    rois[:, 0] = torch.randint(0, 3, (K,), generator=g).to(dtype)
    rois[:, 3:] = rois[:, 1:3] + torch.rand(K, 2, generator=g, dtype=dtype) * 25 + 0.5
    for ph, pw, cm_ in ((2, 3, 2), (1, 1, 4), (3, 3, 1)):
        x = torch.randn(3, cm_ * ph * pw, 17, 23, generator=g, dtype=dtype)
        for scale, sr in ((1.0, 2), (0.5, -1), (0.3, 1)):
            a, ma = oracle.ps_roi_align_forward(x, rois, ph, pw, scale, sr, impl="oracle")
            b, mb = oracle.ps_roi_align_forward(x, rois, ph, pw, scale, sr, impl="ref")
            assert torch.equal(a, b) and torch.equal(ma, mb), (scale, sr, ph, pw)
            go = torch.randn(a.shape, generator=g, dtype=dtype)
            assert torch.equal(oracle.ps_roi_align_backward(go, rois, ma, scale, ph, pw, sr, *x.shape, impl="oracle"),
                               oracle.ps_roi_align_backward(go, rois, mb, scale, ph, pw, sr, *x.shape, impl="ref"))


def test_batched_nms_trick_follows_tensor_max_on_nan():
    """boxes.py:91 takes boxes.max(), which propagates NaN; the restatement must not skip NaN coordinates.
    Checked against the trick written out in torch on top of the oracle's plain nms."""
    g = torch.Generator().manual_seed(4)
    n = 400
    xy = torch.rand(n, 2, generator=g) * 100
    b = torch.cat([xy, xy + torch.rand(n, 2, generator=g) * 40 + 5], 1)
    s = torch.randperm(n, generator=g).float()
    idx = torch.randint(0, 3, (n,), generator=g)
    for where in (None, (0, 0), (123, 2)):
        bb = b.clone()
        if where is not None:
            bb[where] = float("nan")
        offsets = idx.to(bb) * (bb.max() + 1)
        want = oracle.nms(bb + offsets[:, None], s, 0.5)
        assert torch.equal(oracle.batched_nms(bb, s, idx, 0.5, "trick"), want), where
