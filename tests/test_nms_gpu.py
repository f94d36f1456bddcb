"""Parity of the device NMS / batched NMS with the oracle: keep indices must be
bit-exact (torch.equal), through the C-ABI, the host-buffer C-ABI and the
dispatcher.  Mirrors the reference's TestNMS (test/test_ops.py:744-918)."""
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def capi():
    from cpu_vision_b200 import capi as c

    return c


def make_boxes(g, n, span=1000.0, lo=10.0, hi=210.0, distinct=True):
    xy = torch.rand(n, 2, generator=g) * span
    wh = torch.rand(n, 2, generator=g) * (hi - lo) + lo
    scores = torch.randperm(n, generator=g).float() / n if distinct else torch.rand(n, generator=g)
    return torch.cat([xy, xy + wh], 1), scores


def test_golden(golden, capi):
    d = golden("nms.pt")
    b, s = d["boxes"].to(DEV), d["scores"].to(DEV)
    for thr in (0.2, 0.5, 0.7):
        assert torch.equal(capi.nms(b, s, thr).cpu(), d[f"keep_{thr}"])
    assert torch.equal(capi.nms(b, d["scores_ties"].to(DEV), 0.5).cpu(), d["keep_ties"]), "stable tie order"
    i = d["idxs"].to(DEV)
    assert torch.equal(capi.batched_nms(b, s, i, 0.5, capi.BNMS_TRICK).cpu(), d["bn_trick"])
    assert torch.equal(capi.batched_nms(b, s, i, 0.5, capi.BNMS_VANILLA).cpu(), d["bn_vanilla"])
    b2, s2, i2 = d["boxes2"].to(DEV), d["scores2"].to(DEV), d["idxs2"].to(DEV)
    assert torch.equal(capi.batched_nms(b2, s2, i2, 0.5, capi.BNMS_TRICK).cpu(), d["bn2_trick"])
    assert torch.equal(capi.batched_nms(b2, s2, i2, 0.5, capi.BNMS_VANILLA).cpu(), d["bn2_vanilla"])
    k = golden("known_answers.pt")
    assert capi.nms(k["K1_boxes"].to(DEV), k["K1_scores"].to(DEV), 0.2).tolist() == [1]
    assert capi.nms(k["K1_boxes"].half().to(DEV), k["K1_scores"].half().to(DEV), 0.2).tolist() == [1]  # test_nms_float16


@pytest.mark.parametrize("n", [1, 2, 63, 64, 65, 127, 128, 129, 1000, 5000])
@pytest.mark.parametrize("thr", [0.2, 0.5, 0.8])
def test_random_vs_oracle(capi, n, thr):
    for seed in range(3):
        g = torch.Generator().manual_seed(1000 * seed + n)
        b, s = make_boxes(g, n, span=300.0 if n >= 1000 else 100.0)
        want = oracle.nms(b, s, thr)
        got = capi.nms(b.to(DEV), s.to(DEV), thr)
        assert torch.equal(got.cpu(), want), (n, thr, seed)


def test_reference_style_engineered_overlap(capi):
    """test_ops.py:769-795: one box engineered to sit just above the threshold."""
    for thr in (0.2, 0.5, 0.8):
        g = torch.Generator().manual_seed(int(thr * 100))
        n = 1000
        b = torch.rand(n, 4, generator=g) * 100
        b[:, 2:] += b[:, :2]
        b[-1, :] = b[0, :]
        x0, y0, x1, y1 = b[-1].tolist()
        b[-1, 2] += (x1 - x0) * (1 - thr) / thr - 1e-5
        s = torch.rand(n, generator=g)
        assert torch.equal(capi.nms(b.to(DEV), s.to(DEV), thr).cpu(), oracle.nms(b, s, thr))


def test_ties_nan_free_duplicates(capi):
    g = torch.Generator().manual_seed(3)
    b, _ = make_boxes(g, 500, span=80.0)
    b = torch.cat([b, b[:100]])                                   # exact duplicates
    s = torch.cat([torch.full((300,), 0.5), torch.rand(300, generator=g)])  # long tie run
    assert torch.equal(capi.nms(b.to(DEV), s.to(DEV), 0.5).cpu(), oracle.nms(b, s, 0.5))
    z = torch.zeros(10, 4)                                        # degenerate zero-area boxes: 0/0 -> NaN > thr false
    assert torch.equal(capi.nms(z.to(DEV), torch.arange(10.0).to(DEV), 0.5).cpu(), oracle.nms(z, torch.arange(10.0), 0.5))


def test_fp64_and_empty(capi):
    g = torch.Generator().manual_seed(8)
    b, s = make_boxes(g, 700, span=150.0)
    b, s = b.double(), s.double()
    assert torch.equal(capi.nms(b.to(DEV), s.to(DEV), 0.5).cpu(), oracle.nms(b, s, 0.5))
    e = capi.nms(torch.zeros(0, 4, device=DEV), torch.zeros(0, device=DEV), 0.5)
    assert e.shape == (0,) and e.dtype == torch.int64


@pytest.mark.parametrize("n,classes", [(1000, 4), (5000, 80), (3000, 1), (5001, 90)])
def test_batched_vs_oracle(capi, n, classes):
    """cfg2: 5000 boxes x 80 classes, IoU 0.5 -> bit-exact keep indices (both strategies and AUTO)."""
    for seed in range(2):
        g = torch.Generator().manual_seed(1234 + 2 + seed)
        b, s = make_boxes(g, n)
        idx = torch.randint(0, classes, (n,), generator=g)
        bd, sd, idd = b.to(DEV), s.to(DEV), idx.to(DEV)
        assert torch.equal(capi.batched_nms(bd, sd, idd, 0.5, capi.BNMS_TRICK).cpu(),
                           oracle.batched_nms(b, s, idx, 0.5, "trick"))
        assert torch.equal(capi.batched_nms(bd, sd, idd, 0.5, capi.BNMS_VANILLA).cpu(),
                           oracle.batched_nms(b, s, idx, 0.5, "vanilla"))
        assert torch.equal(capi.batched_nms(bd, sd, idd, 0.5, capi.BNMS_AUTO).cpu(),
                           oracle.batched_nms(b, s, idx, 0.5, "auto_cuda"))


def test_batched_nms_implementations_agree(capi):
    """test_ops.py:894-918: vanilla == coordinate trick for N=1000, 4 classes, thr 0.9; empty input."""
    for seed in range(5):
        g = torch.Generator().manual_seed(seed)
        b = torch.rand(1000, 4, generator=g) * 100
        b[:, 2:] += b[:, :2]
        s = torch.randperm(1000, generator=g).float() / 1000
        idx = torch.randint(0, 4, (1000,), generator=g)
        a = capi.batched_nms(b.to(DEV), s.to(DEV), idx.to(DEV), 0.9, capi.BNMS_TRICK)
        c = capi.batched_nms(b.to(DEV), s.to(DEV), idx.to(DEV), 0.9, capi.BNMS_VANILLA)
        assert torch.equal(a, c)
    import cpu_vision_b200 as m

    e = m.batched_nms(torch.zeros(0, 4, device=DEV), torch.zeros(0, device=DEV), torch.zeros(0, dtype=torch.int64, device=DEV), 0.2)
    assert e.numel() == 0 and e.dtype == torch.int64


def test_host_buffer_entry_points(capi):
    g = torch.Generator().manual_seed(21)
    b, s = make_boxes(g, 2000, span=300.0)
    idx = torch.randint(0, 20, (2000,), generator=g)
    assert torch.equal(capi.nms_host(b.pin_memory(), s.pin_memory(), 0.5), oracle.nms(b, s, 0.5))
    assert torch.equal(capi.batched_nms_host(b, s, idx, 0.5), oracle.batched_nms(b, s, idx, 0.5, "auto_cuda"))
    x = torch.randn(2, 64, 30, 40, generator=g)
    rois = torch.tensor([[0, 4.0, 4, 60, 50], [1, 10, 20, 100, 110], [1, 0, 0, 159, 119]])
    out = capi.roi_align_forward_host(x, rois, 7, 7, 0.25, 2, True)
    torch.testing.assert_close(out, oracle.roi_align_forward(x, rois, 7, 7, 0.25, 2, True), rtol=1e-5, atol=1e-6)
    go = torch.randn(3, 64, 7, 7, generator=g)
    gi = capi.roi_align_backward_host(go, rois, 0.25, 7, 7, 2, 64, 30, 40, 2, True)
    torch.testing.assert_close(gi, oracle.roi_align_backward(go, rois, 0.25, 7, 7, 2, 64, 30, 40, 2, True),
                               rtol=1e-4, atol=1e-5)
    capi.host_release()


def test_dispatcher_nms_and_errors():
    import cpu_vision_b200 as m

    g = torch.Generator().manual_seed(33)
    b, s = make_boxes(g, 1500, span=200.0)
    assert torch.equal(torch.ops.torchvision.nms(b.to(DEV), s.to(DEV), 0.5).cpu(), oracle.nms(b, s, 0.5))
    assert torch.equal(m.nms(b.to(DEV), s.to(DEV), 0.7).cpu(), oracle.nms(b, s, 0.7))
    idx = torch.randint(0, 80, (1500,), generator=g)
    assert torch.equal(m.batched_nms(b.to(DEV), s.to(DEV), idx.to(DEV), 0.5).cpu(),
                       oracle.batched_nms(b, s, idx, 0.5, "auto_cuda"))
    # test_ops.py:797-805 input errors
    with pytest.raises(RuntimeError):
        m.nms(torch.rand(4, device=DEV), torch.rand(3, device=DEV), 0.5)
    with pytest.raises(RuntimeError):
        m.nms(torch.rand(3, 5, device=DEV), torch.rand(5, device=DEV), 0.5)
    with pytest.raises(RuntimeError):
        m.nms(torch.rand(3, 4, device=DEV), torch.rand(3, 2, device=DEV), 0.5)
    with pytest.raises(RuntimeError):
        m.nms(torch.rand(3, 4, device=DEV), torch.rand(4, device=DEV), 0.5)


def test_full_size_properties(capi):
    """cfg2 'reading B' scale: properties that hold at any N (no oracle needed):
    no two kept boxes of a class overlap above thr; every dropped box is covered by
    a kept higher-scored box of its class; nms of the kept set keeps everything."""
    g = torch.Generator().manual_seed(77)
    n, classes, thr = 20000, 80, 0.5
    b, s = make_boxes(g, n, span=2000.0)
    idx = torch.randint(0, classes, (n,), generator=g)
    bd, sd, idd = b.to(DEV), s.to(DEV), idx.to(DEV)
    keep = capi.batched_nms(bd, sd, idd, thr, capi.BNMS_VANILLA)
    assert torch.all(sd[keep][:-1] >= sd[keep][1:]), "descending score order"
    kb, kc, ks = bd[keep].double(), idd[keep], sd[keep]

    def iou(a, c):
        lt = torch.maximum(a[:, None, :2], c[None, :, :2])
        rb = torch.minimum(a[:, None, 2:], c[None, :, 2:])
        wh = (rb - lt).clamp(min=0)
        inter = wh[..., 0] * wh[..., 1]
        aa = (a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1])
        ac = (c[:, 2] - c[:, 0]) * (c[:, 3] - c[:, 1])
        return inter / (aa[:, None] + ac[None, :] - inter)

    m = iou(kb, kb)
    same = kc[:, None] == kc[None, :]
    m = torch.where(same, m, torch.zeros_like(m))
    m.fill_diagonal_(0)
    assert (m > thr + 1e-6).sum().item() == 0
    dropped = torch.ones(n, dtype=torch.bool, device=DEV)
    dropped[keep] = False
    di = torch.nonzero(dropped).flatten()
    cover = iou(bd[di].double(), kb)
    ok = (cover > thr - 1e-6) & (idd[di][:, None] == kc[None, :]) & (ks[None, :] > sd[di][:, None])
    assert ok.any(dim=1).all()
    again = capi.batched_nms(bd[keep], sd[keep], idd[keep], thr, capi.BNMS_VANILLA)
    assert torch.equal(again, torch.arange(keep.numel(), device=DEV))


# ---- segmented (per-image) NMS: b200det_nms_segmented ------------------------------------------
def _segmented_oracle(b, s, i, off, thr, valid, mode, top_n):
    """Per image: batched_nms of the surviving boxes, mapped back to positions in the concatenation
    (what the Python loops of rpn.py:268-297 / roi_heads.py:693-723 compute)."""
    out = []
    for k in range(len(off) - 1):
        sel = torch.arange(off[k], off[k + 1])
        if valid is not None:
            sel = sel[valid[sel].bool()]
        if sel.numel() == 0:
            out.append(torch.empty(0, dtype=torch.int64))
            continue
        if i is None:
            keep = oracle.nms(b[sel], s[sel], thr)
        else:
            # the reference's rule (boxes.py:72) sees the boxes AFTER the boxes[keep] compactions
            m = mode if mode != "auto" else ("vanilla" if 4 * sel.numel() > 20000 else "trick")
            keep = oracle.batched_nms(b[sel], s[sel], i[sel], thr, m)
        keep = sel[keep]
        out.append(keep[:top_n] if top_n > 0 else keep)
    return out


@pytest.mark.parametrize("sizes", [[5], [0, 7, 0], [64, 65, 1, 130], [700, 1, 0, 333, 1200], [300] * 70])
@pytest.mark.parametrize("mode", ["plain", "trick", "vanilla"])
def test_segmented_vs_oracle(capi, sizes, mode):
    g = torch.Generator().manual_seed(sum(sizes) + len(mode))
    n = sum(sizes)
    off = [0]
    for z in sizes:
        off.append(off[-1] + z)
    b, s = make_boxes(g, max(n, 1), span=200.0)
    b, s = b[:n], s[:n]
    i = None if mode == "plain" else torch.randint(0, 5, (n,), generator=g)
    for valid, top_n in ((None, 0), (torch.rand(n, generator=g) > 0.3, 0), (None, 50)):
        want = _segmented_oracle(b, s, i, off, 0.5, valid, mode, top_n)
        got = capi.nms_segmented(b.to(DEV), s.to(DEV), None if i is None else i.to(DEV), off, 0.5,
                                 valid=None if valid is None else valid.to(DEV),
                                 mode={"plain": capi.BNMS_AUTO, "trick": capi.BNMS_TRICK, "vanilla": capi.BNMS_VANILLA}[mode],
                                 top_n=top_n)
        assert len(got) == len(want)
        for k, (a, w) in enumerate(zip(got, want)):
            assert torch.equal(a.cpu(), w), (sizes, mode, k, top_n)


def test_segmented_auto_mode_and_negative_coordinates(capi):
    """AUTO picks the strategy per image from its box count (boxes.py:72); with negative coordinates the
    coordinate trick lets boxes of different classes meet, and that reference behaviour is kept."""
    g = torch.Generator().manual_seed(77)
    sizes = [5200, 900]  # vanilla for the first image, trick for the second
    off = [0, 5200, 6100]
    b, s = make_boxes(g, 6100, span=400.0)
    b = b - 150.0
    i = torch.randint(0, 4, (6100,), generator=g)
    want = _segmented_oracle(b, s, i, off, 0.5, None, "auto", 0)
    got = capi.nms_segmented(b.to(DEV), s.to(DEV), i.to(DEV), off, 0.5)
    for a, w in zip(got, want):
        assert torch.equal(a.cpu(), w)


def test_segmented_auto_mode_counts_the_surviving_boxes(capi):
    """A set of 8000 candidates of which the filters keep 3500: the reference compacts first (rpn.py:276-288), so
    batched_nms sees 14 000 elements and takes the coordinate trick; 8000 unfiltered candidates take the vanilla
    strategy.  Negative coordinates make the two strategies differ, so the choice is visible in the result."""
    g = torch.Generator().manual_seed(78)
    n = 8000
    off = [0, n, 2 * n]
    b, s = make_boxes(g, 2 * n, span=120.0, lo=20.0, hi=60.0)
    b = b - 100.0   # offsets idx * (max + 1) no longer separate the classes: the strategies disagree
    i = torch.randint(0, 4, (2 * n,), generator=g)
    valid = torch.zeros(2 * n, dtype=torch.bool)
    valid[torch.randperm(n, generator=g)[:3500]] = True       # first set: 3500 survivors -> trick
    valid[n:] = True                                          # second set: all 8000 -> vanilla
    want = _segmented_oracle(b, s, i, off, 0.5, valid, "auto", 0)
    other = _segmented_oracle(b, s, i, off, 0.5, valid, "vanilla", 0)
    assert not torch.equal(want[0], other[0]), "the case must tell the strategies apart"
    got = capi.nms_segmented(b.to(DEV), s.to(DEV), i.to(DEV), off, 0.5, valid=valid.to(DEV))
    for a, w in zip(got, want):
        assert torch.equal(a.cpu(), w)


@pytest.mark.parametrize("mode", ["trick", "vanilla"])
def test_segmented_class_grouping_keeps_the_reference_result(capi, mode):
    """Class-aware sets are scanned grouped by class with tiles between classes skipped.  The result must stay
    the per-image batched_nms of the oracle when classes are many, ids are negative or large, boxes are filtered,
    top-n cuts the list, and when the coordinate trick does NOT separate the classes (negative, infinite and NaN
    coordinates) -- there the bounding-box test has to keep the tile."""
    g = torch.Generator().manual_seed(91)
    sizes = [1500, 0, 2300, 64, 129]
    n = sum(sizes)
    off = [0]
    for z in sizes:
        off.append(off[-1] + z)
    b, s = make_boxes(g, n, span=150.0)
    i = torch.randint(-3, 80, (n,), generator=g)
    i[::97] = 2_000_000
    valid = torch.rand(n, generator=g) > 0.25
    m = {"trick": capi.BNMS_TRICK, "vanilla": capi.BNMS_VANILLA}[mode]
    for variant in ("clean", "negative", "nonfinite"):
        bb = b.clone()
        if variant == "negative":
            bb = bb - 400.0                       # offsets idx * (max + 1) no longer separate the classes
        if variant == "nonfinite":
            bb[5::301, 0] = float("nan")
            bb[7::311, 3] = float("nan")
            bb[11::401, 2] = float("inf")
            bb[13::409, 0] = float("-inf")
        for v, top_n in ((None, 0), (valid, 0), (valid, 100)):
            want = _segmented_oracle(bb, s, i, off, 0.5, v, mode, top_n)
            got = capi.nms_segmented(bb.to(DEV), s.to(DEV), i.to(DEV), off, 0.5,
                                     valid=None if v is None else v.to(DEV), mode=m, top_n=top_n)
            for k, (a, w) in enumerate(zip(got, want)):
                assert torch.equal(a.cpu(), w), (mode, variant, k, top_n)
        for thr in (-0.5, 0.0):  # a negative threshold makes disjoint boxes of different classes interact
            want = _segmented_oracle(bb, s, i, off, thr, valid, mode, 0)
            got = capi.nms_segmented(bb.to(DEV), s.to(DEV), i.to(DEV), off, thr, valid=valid.to(DEV), mode=m)
            for k, (a, w) in enumerate(zip(got, want)):
                assert torch.equal(a.cpu(), w), (mode, variant, k, thr)
        one = capi.batched_nms(bb[:1500].to(DEV), s[:1500].to(DEV), i[:1500].to(DEV), 0.5, m)
        assert torch.equal(one.cpu(), oracle.batched_nms(bb[:1500], s[:1500], i[:1500], 0.5, mode)), (mode, variant)


def test_segmented_equals_loop_of_batched_nms_full_size(capi):
    """64 images x 4663 RPN candidates (5 levels): one segmented call == 64 batched_nms calls."""
    g = torch.Generator().manual_seed(5)
    per, S = 4663, 64
    b, s = make_boxes(g, per * S, span=1000.0)
    s = torch.cat([torch.randperm(per, generator=g).float() / per for _ in range(S)])  # distinct per image
    lvl = torch.randint(0, 5, (per * S,), generator=g)
    off = [k * per for k in range(S + 1)]
    bd, sd, ld = b.to(DEV), s.to(DEV), lvl.to(DEV)
    got = capi.nms_segmented(bd, sd, ld, off, 0.7, top_n=1000)
    for k in (0, 17, 63):
        sl = slice(off[k], off[k + 1])
        ref = capi.batched_nms(bd[sl], sd[sl], ld[sl], 0.7)[:1000] + off[k]
        assert torch.equal(got[k], ref)
        assert torch.equal(got[k].cpu(), _segmented_oracle(b, s, lvl, off, 0.7, None, "auto", 1000)[k]) if k == 0 else True


def test_iou_on_the_threshold_and_degenerate_boxes(capi):
    """The fp32 mask decides most pairs by the sign of fma(-thr, union, inter) and falls back to the exact
    quotient near the threshold: pairs whose IoU IS the threshold (integer grids: 1/2, 1/3, 1/4 ...), pairs one
    ulp either side of it, inverted / zero-area / huge boxes must all come out as the reference's division does."""
    g = torch.Generator().manual_seed(77)
    n = 1500
    xy = torch.randint(0, 12, (n, 2), generator=g).float()
    wh = torch.randint(1, 5, (n, 2), generator=g).float()
    grid = torch.cat([xy, xy + wh], 1)                       # small integer boxes: many exact rational IoUs
    s = torch.randperm(n, generator=g).float()
    for thr in (0.5, 1.0 / 3.0, 0.25, 0.2, 2.0 / 3.0, 0.75, 1.0, 1e-6, 0.0, -1.0, 1.5):
        assert torch.equal(capi.nms(grid.to(DEV), s.to(DEV), thr).cpu(), oracle.nms(grid, s, thr)), thr
    # one ulp around: widen one side of a duplicate so that IoU = w / (w + d) sweeps through 0.5
    base = torch.tensor([0.0, 0.0, 1.0, 1.0])
    rows, x = [base], torch.tensor(2.0)
    for _ in range(40):
        x = torch.nextafter(x, torch.tensor(0.0))
        rows.append(torch.stack([base[0], base[1], x, base[3]]))
    x = torch.tensor(2.0)
    for _ in range(40):
        x = torch.nextafter(x, torch.tensor(4.0))
        rows.append(torch.stack([base[0], base[1], x, base[3]]))
    near = torch.stack(rows)
    sn = torch.arange(len(rows), 0, -1).float()
    for thr in (0.5, float(torch.nextafter(torch.tensor(0.5), torch.tensor(1.0))),
                float(torch.nextafter(torch.tensor(0.5), torch.tensor(0.0)))):
        assert torch.equal(capi.nms(near.to(DEV), sn.to(DEV), thr).cpu(), oracle.nms(near, sn, thr)), thr
    # inverted (negative area), zero-area, huge and tiny boxes mixed into a normal set
    b, s2 = make_boxes(g, 600, span=60.0)
    b[::7, 2:] = b[::7, :2] - 3.0                            # x2 < x1, y2 < y1
    b[3::11, 2:] = b[3::11, :2]                              # zero area
    b[5::13] *= 1e18                                         # areas overflow to inf
    b[6::17] *= 1e-22                                        # areas underflow
    for thr in (0.3, 0.5):
        assert torch.equal(capi.nms(b.to(DEV), s2.to(DEV), thr).cpu(), oracle.nms(b, s2, thr)), thr
    idx = torch.randint(0, 4, (600,), generator=g)
    for mode, name in ((capi.BNMS_TRICK, "trick"), (capi.BNMS_VANILLA, "vanilla")):
        got = capi.batched_nms(grid[:600].to(DEV), s[:600].to(DEV), idx.to(DEV), 0.5, mode)
        assert torch.equal(got.cpu(), oracle.batched_nms(grid[:600], s[:600], idx, 0.5, name)), name


@pytest.mark.parametrize("n", [7, 4096, 8191, 8192, 8193, 13312, 13313])
def test_sort_front_ends_agree_with_the_oracle_on_special_scores(capi, n):
    """Sets of up to 8192 boxes take the single-CTA sort, larger ones the device-wide radix sort: signed zeros,
    infinities, long tie runs and 16-bit scores must rank the same way in both, i.e. like the oracle's stable
    descending sort."""
    g = torch.Generator().manual_seed(n)
    b, _ = make_boxes(g, n, span=400.0)
    s = torch.rand(n, generator=g)
    s[::5] = 0.0
    s[1::5] = -0.0
    s[2::50] = float("inf")
    s[3::50] = float("-inf")
    s[4::9] = 0.25
    s = s * torch.where(torch.rand(n, generator=g) < 0.3, -1.0, 1.0)
    assert torch.equal(capi.nms(b.to(DEV), s.to(DEV), 0.5).cpu(), oracle.nms(b, s, 0.5))
    idx = torch.randint(0, 6, (n,), generator=g)
    for mode, name in ((capi.BNMS_TRICK, "trick"), (capi.BNMS_VANILLA, "vanilla")):
        got = capi.batched_nms(b.to(DEV), s.to(DEV), idx.to(DEV), 0.5, mode)
        assert torch.equal(got.cpu(), oracle.batched_nms(b, s, idx, 0.5, name)), name
    bh, sh = b.half(), s.half()
    assert torch.equal(capi.nms(bh.to(DEV), sh.to(DEV), 0.5).cpu(), oracle.nms(bh.float(), sh.float(), 0.5))


@pytest.mark.parametrize("n", [1, 3, 64, 65, 257, 5000, 8192, 10000, 13312])
def test_front_ends_rank_sort_and_block_sort(capi, n, monkeypatch):
    """The rank sort over all SMs (default for sets of up to 8192 boxes) and the one-CTA block radix sort
    (B200DET_NMS_FRONT=block, read per call) must give the same keep list, which is the oracle's: tie runs,
    signed zeros, infinities and NaN scores included (NaN sorts first, like torch's descending sort)."""
    g = torch.Generator().manual_seed(100 + n)
    b, _ = make_boxes(g, n, span=300.0)
    s = torch.rand(n, generator=g)
    s[::4] = 0.5
    s[1::6] = 0.0
    s[2::6] = -0.0
    s[3::40] = float("inf")
    s[5::40] = float("-inf")
    s = s * torch.where(torch.rand(n, generator=g) < 0.3, -1.0, 1.0)
    idx = torch.randint(0, 5, (n,), generator=g)
    want = oracle.nms(b, s, 0.5)
    wants = {name: oracle.batched_nms(b, s, idx, 0.5, name) for name in ("trick", "vanilla")}
    sn = s.clone()
    sn[7::31] = float("nan")
    out = {}
    for front in ("rank", "block"):
        monkeypatch.setenv("B200DET_NMS_FRONT", front)
        assert torch.equal(capi.nms(b.to(DEV), s.to(DEV), 0.5).cpu(), want), front
        for mode, name in ((capi.BNMS_TRICK, "trick"), (capi.BNMS_VANILLA, "vanilla")):
            got = capi.batched_nms(b.to(DEV), s.to(DEV), idx.to(DEV), 0.5, mode)
            assert torch.equal(got.cpu(), wants[name]), (front, name)
        out[front] = capi.nms(b.to(DEV), sn.to(DEV), 0.5).cpu()
        bh, sh = b.half(), s.half()
        assert torch.equal(capi.nms(bh.to(DEV), sh.to(DEV), 0.5).cpu(), oracle.nms(bh.float(), sh.float(), 0.5)), front
    assert torch.equal(out["rank"], out["block"])


# ---- the fixed-point resolve kernel (all SMs) and its fall-back to the serial scan --------------------------
def _chain(n, kind, w=10.0, h=10.0):
    """Boxes on a line, scores descending with the index.
    deep:     box i shifted by 2 i: IoU(i, i + 1) = 8/12 > 0.5 and IoU(i, i + 2) = 6/14 < 0.5, so the greedy result
              alternates kept / removed down the line -- a suppression chain as deep as the set.
    clusters: groups of 5 boxes one unit apart (IoU 0.82, 0.67, 0.54, 0.43 with the first), groups far apart: box 0
              of a group removes 1-3, box 4 survives only because they are gone -- depth 2 everywhere."""
    i = torch.arange(n, dtype=torch.float32)
    x = i * 2.0 if kind == "deep" else torch.floor(i / 5) * 100.0 + torch.remainder(i, 5) * 1.0
    boxes = torch.stack([x, torch.zeros(n), x + w, torch.full((n,), h)], 1)
    scores = torch.linspace(1.0, 0.5, n)
    return boxes, scores


@pytest.mark.parametrize("resolve", ["1", "0"])   # B200DET_NMS_RESOLVE: fixed point on all SMs first / serial scan alone
@pytest.mark.parametrize("kind", ["deep", "clusters"])
@pytest.mark.parametrize("n", [9, 130, 3000, 8192, 13312])
def test_suppression_chains_resolve_kernel_and_fallback(capi, n, kind, resolve, monkeypatch):
    """Deep chains need ~n/2 rounds of the fixed-point iteration, far beyond the resolve kernel's limit: the serial
    scan must take over.  Clusters are finished by the resolve kernel itself.  Same keep indices as the oracle
    either way, also when the chains sit among random boxes of several classes (both batched strategies)."""
    monkeypatch.setenv("B200DET_NMS_RESOLVE", resolve)
    boxes, scores = _chain(n, kind)
    keep = capi.nms(boxes.to(DEV), scores.to(DEV), 0.5)
    want = oracle.nms(boxes, scores, 0.5)
    assert torch.equal(keep.cpu(), want)
    assert want.numel() == ((n + 1) // 2 if kind == "deep" else 2 * (n // 5) + min(n % 5, 1))
    g = torch.Generator().manual_seed(n)
    m = min(n, 2000)
    xy = torch.rand(m, 2, generator=g) * 300
    rb = torch.cat([xy, xy + torch.rand(m, 2, generator=g) * 60 + 5], 1)
    k = min(n, 8192 - m)
    ab, asc = torch.cat([boxes[:k], rb]), torch.cat([scores[:k], torch.rand(m, generator=g)])
    idxs = torch.randint(0, 3, (ab.shape[0],), generator=g)
    for mode in (capi.BNMS_TRICK, capi.BNMS_VANILLA):
        keep = capi.batched_nms(ab.to(DEV), asc.to(DEV), idxs.to(DEV), 0.5, mode)
        assert torch.equal(keep.cpu(), oracle.batched_nms(ab, asc, idxs, 0.5, "trick" if mode == capi.BNMS_TRICK else "vanilla"))
