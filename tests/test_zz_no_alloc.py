"""b200det_nms_static: the allocation-free NMS entry point (scratch in a fixed array of the library's device
image, include/b200_detection_ops.h).  It has no counterpart in the reference; the check is the existing entry
point b200det_nms on the same inputs, bit-exact like every NMS test here.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def capi():
    from cpu_vision_b200 import capi as m

    m.lib()
    return m


def _boxes(g, n, span):
    xy = torch.rand(n, 2, generator=g) * span
    wh = torch.rand(n, 2, generator=g) * 200.0 + 10.0
    s = torch.rand(n, generator=g)
    s[::7] = 0.5  # tie runs: the stable order decides
    return torch.cat([xy, xy + wh], 1), s


def _static(capi, b, s, thr, keep, cnt):
    """One call through the C-ABI with every buffer already allocated; reads the result back afterwards."""
    capi.nms_static(b, s, thr, keep, cnt)
    return keep[: int(cnt.item())].clone()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_two_calls_in_a_row_match_b200det_nms(capi, dtype):
    g = torch.Generator().manual_seed(77)
    # a large set first, then a small one on the same (now dirty) scratch, then the limit
    sets = [_boxes(g, 5000, 900.0), _boxes(g, 300, 150.0), _boxes(g, capi.NMS_STATIC_MAX_BOXES, 1500.0)]
    keep = torch.empty(capi.NMS_STATIC_MAX_BOXES, dtype=torch.int64, device=DEV)
    cnt = torch.empty(1, dtype=torch.int64, device=DEV)
    for b, s in sets:
        bd, sd = b.to(DEV, dtype).contiguous(), s.to(DEV, dtype).contiguous()
        want = capi.nms(bd, sd, 0.5)
        assert 0 < want.numel() < bd.shape[0]
        got1 = _static(capi, bd, sd, 0.5, keep, cnt)
        got2 = _static(capi, bd, sd, 0.5, keep, cnt)
        assert torch.equal(got1, want)
        assert torch.equal(got2, want)


def test_back_to_back_without_a_sync_between_the_calls(capi):
    """Two different sets queued on one stream: the second call reuses the scratch as soon as the first one's
    kernels are done (stream order), each result goes to its own output."""
    g = torch.Generator().manual_seed(78)
    (b1, s1), (b2, s2) = _boxes(g, 4000, 700.0), _boxes(g, 1234, 300.0)
    b1, s1, b2, s2 = (t.to(DEV) for t in (b1, s1, b2, s2))
    k1, k2 = (torch.empty(4000, dtype=torch.int64, device=DEV) for _ in range(2))
    c1, c2 = (torch.empty(1, dtype=torch.int64, device=DEV) for _ in range(2))
    capi.nms_static(b1, s1, 0.3, k1, c1)
    capi.nms_static(b2, s2, 0.7, k2, c2)
    torch.cuda.synchronize()
    assert torch.equal(k1[: int(c1)], capi.nms(b1, s1, 0.3))
    assert torch.equal(k2[: int(c2)], capi.nms(b2, s2, 0.7))


def test_the_call_allocates_nothing(capi):
    g = torch.Generator().manual_seed(79)
    b, s = _boxes(g, 2048, 500.0)
    b, s = b.to(DEV), s.to(DEV)
    keep = torch.empty(2048, dtype=torch.int64, device=DEV)
    cnt = torch.empty(1, dtype=torch.int64, device=DEV)
    capi.nms_static(b, s, 0.5, keep, cnt)  # set-up call: the symbol address is looked up once per device
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    torch_bytes0 = torch.cuda.memory_allocated()
    for _ in range(5):
        capi.nms_static(b, s, 0.5, keep, cnt)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == torch_bytes0
    assert free0 - torch.cuda.mem_get_info()[0] < (1 << 20)  # no cudaMalloc behind torch's back either
    assert torch.equal(keep[: int(cnt)], capi.nms(b, s, 0.5))


def test_empty_set_and_sizes_beyond_the_limit(capi):
    lib = capi.lib()
    cnt = torch.full((1,), -1, dtype=torch.int64, device=DEV)
    keep = torch.empty(1, dtype=torch.int64, device=DEV)
    e = torch.empty(0, 4, device=DEV)
    capi.nms_static(e, torch.empty(0, device=DEV), 0.5, keep, cnt)
    assert int(cnt) == 0
    n = capi.NMS_STATIC_MAX_BOXES + 1
    b = torch.zeros(n, 4, device=DEV)
    s = torch.zeros(n, device=DEV)
    k = torch.empty(n, dtype=torch.int64, device=DEV)
    with pytest.raises(capi.B200DetError, match="exceeds the built-in scratch"):
        capi.nms_static(b, s, 0.5, k, cnt)
    rc = lib.b200det_nms_static(b.data_ptr(), s.data_ptr(), 0, C.c_int64(-1), 0.5, k.data_ptr(), cnt.data_ptr(), None)
    assert rc == -1 and b"negative N" in lib.b200det_last_error()
    assert int(cnt) == 0  # the refused calls wrote nothing
