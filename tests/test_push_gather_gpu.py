"""b200det_msroi_align_forward_push on ONE GPU: the "peer" buffers are further buffers of the same device, so the
push variant of the forward kernel (every pooled bin stored to the output and to the same row behind each peer
pointer) is exercised wherever the GPU tests run; across real peers over NVLink it is tests/test_multigpu_gpu.py.
Bar: the pushed copies are BITWISE the plain forward's output, which is checked against the oracle."""
import pytest
import torch

import oracle
from test_roi_align_gpu import DEV, close_fwd, rand_rois

pytestmark = pytest.mark.gpu

SHAPES = [(50, 68), (25, 34)]
SCALES = [0.0625, 0.03125]


@pytest.fixture(scope="module")
def capi():
    from cpu_vision_b200 import capi as c

    return c


def _case(dtype, C, K, seed=5):
    g = torch.Generator().manual_seed(seed)
    feats = [torch.randn(2, C, h, w, generator=g).to(dtype) for h, w in SHAPES]
    rois = rand_rois(g, K, 2, 1088.0, 800.0, lo=16.0, hi=800.0)
    return feats, rois


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("C,ph,sr,peers", [(256, 7, 2, 1), (256, 14, 2, 3), (128, 7, 1, 7)])
def test_push_copies_equal_the_plain_forward(capi, dtype, C, ph, sr, peers):
    feats, rois = _case(dtype, C, 96)
    fd = [f.to(DEV).contiguous(memory_format=torch.channels_last) for f in feats]
    rd = rois.to(DEV)
    want, lv = capi.msroi_align_forward(fd, rd, SCALES, ph, ph, sr, False, k_min=4, k_max=5)
    K = rois.shape[0]
    # one "gather buffer" per fake peer, this rank's slice in the middle of it
    bufs = [torch.full((3 * K, C, ph, ph), float("nan"), dtype=dtype, device=DEV).contiguous(memory_format=torch.channels_last)
            for _ in range(peers)]
    mine = torch.empty_like(want)
    ptrs = [b[K:2 * K].data_ptr() for b in bufs]
    out, lv2 = capi.msroi_align_forward_push(fd, rd, SCALES, ph, ph, sr, mine, ptrs, k_min=4, k_max=5)
    torch.cuda.synchronize()
    assert torch.equal(lv, lv2)
    assert torch.equal(out, want)
    for b in bufs:
        assert torch.equal(b[K:2 * K], want)
        assert torch.isnan(b[:K]).all() and torch.isnan(b[2 * K:]).all()   # nothing outside the slice is touched
    if dtype == torch.float32:
        levels = oracle.fpn_level_map(rois, 4, 5)
        close_fwd(out, oracle.msroi_align_forward(feats, rois, levels, SCALES, ph, ph, sr, False))


def test_push_without_peers_and_empty_rois(capi):
    feats, rois = _case(torch.float32, 256, 32)
    fd = [f.to(DEV).contiguous(memory_format=torch.channels_last) for f in feats]
    want, _ = capi.msroi_align_forward(fd, rois.to(DEV), SCALES, 7, 7, 2, False, k_min=4, k_max=5)
    out, _ = capi.msroi_align_forward_push(fd, rois.to(DEV), SCALES, 7, 7, 2, torch.empty_like(want), [], k_min=4, k_max=5)
    assert torch.equal(out, want)
    empty = torch.empty(0, 256, 7, 7, device=DEV).contiguous(memory_format=torch.channels_last)
    out, _ = capi.msroi_align_forward_push(fd, rois[:0].to(DEV), SCALES, 7, 7, 2, empty, [empty.data_ptr() or 16],
                                           k_min=4, k_max=5)
    assert out.shape[0] == 0


def test_push_refuses_what_the_fast_path_does_not_take(capi):
    """Adaptive sampling goes to the any-grid kernels, which do not push: a clear error, not a silent local-only run."""
    feats, rois = _case(torch.float32, 256, 16)
    fd = [f.to(DEV).contiguous(memory_format=torch.channels_last) for f in feats]
    mine = torch.empty(16, 256, 7, 7, device=DEV).contiguous(memory_format=torch.channels_last)
    other = torch.empty_like(mine)
    with pytest.raises(RuntimeError, match="only the channels_last fast path pushes"):
        capi.msroi_align_forward_push(fd, rois.to(DEV), SCALES, 7, 7, 0, mine, [other.data_ptr()], k_min=4, k_max=5)
    with pytest.raises(RuntimeError, match="at most 7 peer outputs"):
        capi.msroi_align_forward_push(fd, rois.to(DEV), SCALES, 7, 7, 2, mine, [other.data_ptr()] * 8, k_min=4, k_max=5)


def test_peer_memory_helpers_round_trip_in_one_process(capi):
    """alloc / export / free of the C-ABI (open needs a second process: tests/test_multigpu_gpu.py)."""
    p = capi.peer_alloc(1 << 20, DEV)
    assert p % 256 == 0
    h = capi.peer_export(p, DEV)
    assert isinstance(h, bytes) and len(h) == 64 and any(h)
    capi.peer_free(p, DEV)
