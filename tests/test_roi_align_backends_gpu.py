"""Element-wise parity of EVERY backward kernel at BASELINE sizes, and of the fast-path shapes the small cases
of test_roi_align_gpu.py do not reach.

Backends (cpu_vision_b200/csrc/c_api.cu backward_impl):
  scatter        fused zero fill + vector-red scatter (default); B200DET_BWD_MODE=0 / 1 picks the pooled-row
                 pair sweep / the feature-row gather
  strip          write-once strip kernel for every level (B200DET_BWD_STRIP=1)
  hybrid         strips for the stride-4 level, scatter for the rest (B200DET_STRIP_MIN_MB)
  deterministic  the strip kernel through the B200DET_BWD_DETERMINISTIC flag
Oracle: the C restatement in fp64 (order-free truth); bar = atomic-reordering tolerance 1e-4 of the gradient's
scale, as in test_roi_align_gpu.py.  The persistent schedulers (zero chunks claimed from a counter, "scatter only
after every chunk is reported done", strips taken from a global queue) only have many chunks / strips in flight at
these sizes, and a red that lands before its zero chunk loses ONE contribution: only an element-wise check sees it.
"""
import os

import pytest
import torch

import oracle
from test_roi_align_gpu import DEV, close_bwd, close_fwd, rand_rois

pytestmark = pytest.mark.gpu

SHAPES = [(200, 272), (100, 136), (50, 68), (25, 34)]
SCALES = [0.25, 0.125, 0.0625, 0.03125]

BACKENDS = {
    "scatter": {},
    "scatter_pair": {"B200DET_BWD_MODE": "0"},
    "scatter_rows": {"B200DET_BWD_MODE": "1"},
    "strip": {"B200DET_BWD_STRIP": "1"},
    "strip_ch4": {"B200DET_BWD_STRIP": "1", "B200DET_STRIP_CH": "4"},   # 128-channel strips of 24 columns
    "hybrid": {"B200DET_STRIP_MIN_MB": "64"},
    "deterministic": {},
}


@pytest.fixture(scope="module")
def capi():
    from cpu_vision_b200 import capi as c

    return c


class backend:
    """Sets the environment switches of one backend for the duration of a call (they are read per call)."""

    def __init__(self, name):
        self.env = BACKENDS[name]
        self.deterministic = name == "deterministic"

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in ("B200DET_BWD_MODE", "B200DET_BWD_STRIP", "B200DET_STRIP_MIN_MB",
                                                   "B200DET_STRIP_CH")}
        for k in self.old:
            os.environ.pop(k, None)
        os.environ.update(self.env)
        return self

    def __exit__(self, *a):
        for k, v in self.old.items():
            os.environ.pop(k, None)
            if v is not None:
                os.environ[k] = v


# ---------------------------------------------------------------- fast-path shapes (SR = 1, non-square, limits)
FAST_CASES = [
    # (N, C, H, W, K, PH, PW, scale, sr, aligned)
    (2, 256, 50, 68, 64, 7, 7, 0.0625, 1, True),      # SR = 1 instantiations
    (2, 256, 50, 68, 64, 7, 7, 0.0625, 1, False),
    (2, 128, 40, 40, 32, 14, 14, 0.25, 1, False),     # SR = 1, 4 channels per lane
    (2, 256, 50, 68, 48, 5, 16, 0.0625, 2, False),    # non-square bins, PW * sr = 32 (the limit)
    (2, 128, 40, 40, 24, 32, 16, 0.25, 2, True),      # PH = 32 (the limit), PW * sr = 32
    (2, 256, 64, 80, 40, 16, 3, 0.125, 1, True),      # tall bins, PH * sr = 16: row gather by default
]


@pytest.mark.parametrize("name", ["scatter_pair", "scatter_rows", "strip", "strip_ch4", "deterministic"])
@pytest.mark.parametrize("case", FAST_CASES)
def test_fast_path_shapes_fp32(capi, case, name):
    N, C, H, W, K, PH, PW, scale, sr, aligned = case
    g = torch.Generator().manual_seed(hash(case) % (2 ** 31))
    x = torch.randn(N, C, H, W, generator=g)
    rois = rand_rois(g, K, N, W / scale, H / scale, lo=1.0 / scale)
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    out = capi.roi_align_forward(xd, rois.to(DEV), PH, PW, scale, sr, aligned)
    close_fwd(out, oracle.roi_align_forward(x, rois, PH, PW, scale, sr, aligned))
    go = torch.randn(K, C, PH, PW, generator=g)
    want64 = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PW, N, C, H, W, sr, aligned)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    with backend(name) as b:
        gi = capi.roi_align_backward(god, rois.to(DEV), scale, PH, PW, N, C, H, W, sr, aligned, layout=capi.NHWC,
                                     deterministic=b.deterministic)
    close_bwd(gi, want64)


@pytest.mark.parametrize("dtype,btol", [(torch.bfloat16, 5e-2), (torch.float16, 2e-2)])
@pytest.mark.parametrize("name", ["scatter_pair", "scatter_rows", "strip"])
@pytest.mark.parametrize("case", FAST_CASES[:4])
def test_fast_path_shapes_16bit(capi, case, name, dtype, btol):
    N, C, H, W, K, PH, PW, scale, sr, aligned = case
    g = torch.Generator().manual_seed(hash(case) % (2 ** 31) + 1)
    x = torch.randn(N, C, H, W, generator=g).to(dtype)
    rois = rand_rois(g, K, N, W / scale, H / scale, lo=1.0 / scale)
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    out = capi.roi_align_forward(xd, rois.to(DEV), PH, PW, scale, sr, aligned)
    want = oracle.roi_align_forward(x.float(), rois, PH, PW, scale, sr, aligned).to(dtype).float()
    torch.testing.assert_close(out.float().cpu().contiguous(), want, rtol=1e-2, atol=1e-2)
    go = torch.randn(K, C, PH, PW, generator=g).to(dtype)
    want64 = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PW, N, C, H, W, sr, aligned)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    with backend(name):
        gi = capi.roi_align_backward(god, rois.to(DEV), scale, PH, PW, N, C, H, W, sr, aligned, layout=capi.NHWC)
    # the strips accumulate in fp32 and round once: far inside the bar of the packed 16-bit reds
    close_bwd(gi.float(), want64, rtol=1e-2 if name == "strip" else btol)


# ---------------------------------------------------------------- BASELINE sizes, element-wise
def _pyramid_case(seed, images, per_img, ph, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    rois = rand_rois(g, images * per_img, images, 1088.0, 800.0, lo=16.0, hi=800.0)
    levels = oracle.fpn_level_map(rois, 2, 5)
    go = torch.randn(images * per_img, 256, ph, ph, generator=g).to(dtype)
    shapes = [(images, 256, h, w) for h, w in SHAPES]
    # fp64 truth level by level (oracle.msroi_align_backward itself computes in fp32)
    want = []
    for l, (h, w) in enumerate(SHAPES):
        sel = (levels == l).nonzero().flatten()
        want.append(oracle.roi_align_backward(go[sel].double(), rois[sel].double(), SCALES[l], ph, ph, images, 256, h, w,
                                              2, False))
    return rois, levels, go, shapes, want


@pytest.fixture(scope="module")
def cfg3_case():
    """cfg3 on whole images: every one of the 512 RoIs of 2 images, all 4 levels, 14x14 bins."""
    return _pyramid_case(1234 + 3, 2, 512, 14)


@pytest.fixture(scope="module")
def cfg4_case():
    """One slice of cfg4: 4 images x 2000 RoIs, 7x7 bins (the row gather / MODE 1 by default)."""
    return _pyramid_case(1234 + 4, 4, 2000, 7)


def _run_pyramid(capi, case, name, batch_pad=0):
    rois, levels, go, shapes, want = case
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    with backend(name) as b:
        gis = capi.msroi_align_backward(god, rois.to(DEV), levels.to(DEV).int(), SCALES, shapes, go.shape[2],
                                        go.shape[3], 2, False, layout=capi.NHWC, deterministic=b.deterministic)
    for gi, w in zip(gis, want):
        close_bwd(gi, w)


@pytest.mark.parametrize("name", list(BACKENDS))
def test_cfg3_backward_elementwise(capi, cfg3_case, name):
    _run_pyramid(capi, cfg3_case, name)


@pytest.mark.parametrize("name", ["scatter", "scatter_pair", "strip", "hybrid"])
def test_cfg4_slice_backward_elementwise(capi, cfg4_case, name):
    _run_pyramid(capi, cfg4_case, name)


@pytest.mark.parametrize("name", ["scatter", "scatter_pair", "strip"])
def test_cfg1_backward_elementwise(capi, name):
    """cfg1 whole: 2 x 256 x 200 x 272, 1000 RoIs per image, 7x7, sr 2, aligned=True, one level of 111 MB."""
    g = torch.Generator().manual_seed(1234 + 1)
    N, C, H, W, K = 2, 256, 200, 272, 2000
    rois = rand_rois(g, K, N, 1088.0, 800.0, lo=16.0, hi=512.0)
    go = torch.randn(K, C, 7, 7, generator=g)
    want64 = oracle.roi_align_backward(go.double(), rois.double(), 0.25, 7, 7, N, C, H, W, 2, True)
    god = go.to(DEV).contiguous(memory_format=torch.channels_last)
    with backend(name):
        gi = capi.roi_align_backward(god, rois.to(DEV), 0.25, 7, 7, N, C, H, W, 2, True, layout=capi.NHWC)
    close_bwd(gi, want64)


def test_cfg3_full_batch_backends_agree(capi):
    """All 16 images x 512 RoIs (the bench workload): the strip kernel (fixed summation order, no zero fill, no
    atomics) and the scatter (fused zero fill, reds in any order) are independent implementations; they must
    agree element-wise within the reordering tolerance on every level, and the strip result must be bitwise
    reproducible."""
    g = torch.Generator().manual_seed(99)
    B, PH = 16, 14
    rois = rand_rois(g, B * 512, B, 1088.0, 800.0, lo=16.0, hi=800.0).to(DEV)
    levels = capi.fpn_level_map(rois, 2, 5)
    go = torch.randn(B * 512, 256, PH, PH, generator=g, dtype=torch.float32).to(DEV).contiguous(memory_format=torch.channels_last)
    shapes = [(B, 256, h, w) for h, w in SHAPES]
    run = lambda det: capi.msroi_align_backward(go, rois, levels, SCALES, shapes, PH, PH, 2, False, layout=capi.NHWC,
                                                deterministic=det)
    with backend("scatter"):
        a = run(False)
    with backend("deterministic"):
        b, c = run(True), run(True)
    for x, y, z in zip(a, b, c):
        assert torch.equal(y, z)
        scale = y.abs().max().item()
        assert (x - y).abs().max().item() <= 1e-4 * scale, ((x - y).abs().max().item(), scale)
