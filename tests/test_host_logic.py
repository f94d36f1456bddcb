"""No-GPU checks of the host-side mirror of the reference's Python front-end and
of the shard-by-image harness (world_size-2 gloo)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_setup_scales_matches_reference_k7(golden):
    from cpu_vision_b200.poolers import _setup_scales

    feats = [torch.zeros(1, 1, 800 // (4 << i), 1088 // (4 << i)) for i in range(4)]
    scales, mapper = _setup_scales(feats, [(800, 1088)], 224, 4)
    d = golden("msroi_align.pt")
    assert scales == d["K7_scales"] == [0.25, 0.125, 0.0625, 0.03125]
    assert (mapper.k_min, mapper.k_max) == (d["K7_kmin"], d["K7_kmax"]) == (2, 5)
    with pytest.raises(ValueError):
        _setup_scales(feats, [], 224, 4)


def test_roi_format_helpers():
    from cpu_vision_b200._utils import check_roi_boxes_shape, convert_boxes_to_roi_format
    from cpu_vision_b200.poolers import _convert_to_roi_format, _filter_input

    a, b = torch.rand(3, 4), torch.rand(2, 4)
    r = convert_boxes_to_roi_format([a, b])
    assert r.shape == (5, 5) and r[:, 0].tolist() == [0, 0, 0, 1, 1] and torch.equal(r[3:, 1:], b)
    assert torch.equal(r, _convert_to_roi_format([a, b]))
    check_roi_boxes_shape([a, b])
    check_roi_boxes_shape(r)
    with pytest.raises(AssertionError):
        check_roi_boxes_shape(torch.rand(3, 4))
    with pytest.raises(AssertionError):
        check_roi_boxes_shape([torch.rand(3, 5)])
    x = {"0": 1, "1": 2, "pool": 3}
    assert _filter_input(x, ["0", "1"]) == [1, 2]


def test_module_reprs_match_reference_format():
    from cpu_vision_b200 import MultiScaleRoIAlign, RoIAlign

    # formats asserted by the reference's test_ops.py::TestMultiScaleRoIAlign::test_msroialign_repr
    m = MultiScaleRoIAlign(["0"], (7, 7), 2)
    assert repr(m) == "MultiScaleRoIAlign(featmap_names=['0'], output_size=(7, 7), sampling_ratio=2)"
    assert repr(RoIAlign(7, 0.25, 2, True)) == "RoIAlign(output_size=7, spatial_scale=0.25, sampling_ratio=2, aligned=True)"


def test_image_block_and_shard_rois():
    from cpu_vision_b200.sharding import image_block, shard_rois

    for n, w in ((64, 8), (10, 4), (3, 4), (16, 1)):
        blocks = [image_block(n, r, w) for r in range(w)]
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        assert all(blocks[i][1] == blocks[i + 1][0] for i in range(w - 1))
        sizes = [hi - lo for lo, hi in blocks]
        assert max(sizes) - min(sizes) <= 1
    rois = torch.tensor([[0.0, 1, 1, 2, 2], [3, 0, 0, 1, 1], [1, 5, 5, 6, 6], [2, 0, 0, 9, 9]])
    loc, sel = shard_rois(rois, 4, 1, 2)
    assert sel.tolist() == [1, 3] and loc[:, 0].tolist() == [1.0, 0.0]
    assert torch.equal(loc[:, 1:], rois[sel, 1:])


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cpu_vision_b200 import sharding

    try:
        # equal rows: in-place gather target
        full, mine = sharding.gather_buffer(3, (2, 2), torch.float32, "cpu", rank, world)
        mine.fill_(float(rank + 1))
        dist.all_gather_into_tensor(full, mine.clone())
        ok = all(bool((full[r * 3:(r + 1) * 3] == r + 1).all()) for r in range(world))
        g = sharding.all_gather_rows(torch.full((3, 2), float(rank)))
        ok &= g.shape == (3 * world, 2) and g[:3].eq(0).all().item() and g[3:6].eq(1).all().item()
        # ragged rows
        g = sharding.all_gather_rows(torch.full((rank + 1, 2), float(rank)))
        ok &= g.shape == (3, 2) and g[:, 0].tolist() == [0.0, 1.0, 1.0]
        # kept-index lists of different length
        keep = torch.arange(10, dtype=torch.int64) + 100 * rank
        cnt = torch.tensor([2 + 3 * rank])
        lists = sharding.all_gather_keep(keep, cnt, capacity=8)
        ok &= [l.tolist() for l in lists] == [[0, 1], [100, 101, 102, 103, 104]]
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gather():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]


def test_dropin_enable_patches_and_falls_through_on_cpu():
    """cpu_vision_b200.dropin under the reference's package (baseline/_ref, a process of its own): enable() replaces
    the three entry points, CPU tensors fall through to the reference's own code (no CPU fallback of ours exists),
    disable() puts the reference's functions back."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = os.path.join(root, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref, "torchvision", "_C.so")):
        import pytest

        pytest.skip("baseline/_ref not built (python baseline/build_ref.py)")
    code = r'''
import torch, torchvision
import cpu_vision_b200
from cpu_vision_b200 import dropin
from torchvision.models.detection import rpn, roi_heads
from torchvision.ops import poolers, MultiScaleRoIAlign
assert "_ref" in torchvision.__file__
orig = (rpn.RegionProposalNetwork.filter_proposals, roi_heads.RoIHeads.postprocess_detections, poolers._multiscale_roi_align)
torch.manual_seed(0)
feats = {"0": torch.randn(2, 8, 32, 32), "1": torch.randn(2, 8, 16, 16)}
boxes = [torch.tensor([[4.0, 4.0, 60.0, 50.0], [10.0, 20.0, 120.0, 100.0]]), torch.tensor([[0.0, 0.0, 30.0, 30.0]])]
pool = MultiScaleRoIAlign(["0", "1"], 3, 2)
want = pool(feats, boxes, [(128, 128), (128, 128)])
names = dropin.enable()
assert len(names) == 3 and dropin.enable() == names                      # idempotent
now = (rpn.RegionProposalNetwork.filter_proposals, roi_heads.RoIHeads.postprocess_detections, poolers._multiscale_roi_align)
assert all(a is not b for a, b in zip(orig, now))
got = pool(feats, boxes, [(128, 128), (128, 128)])                        # CPU tensors: the reference's own loop
assert torch.equal(got, want)
dropin.disable()
back = (rpn.RegionProposalNetwork.filter_proposals, roi_heads.RoIHeads.postprocess_detections, poolers._multiscale_roi_align)
assert all(a is b for a, b in zip(orig, back))
print("dropin cpu ok")
'''
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([ref, root, env.get("PYTHONPATH", "")])
    p = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "dropin cpu ok" in p.stdout, (p.stdout[-2000:], p.stderr[-2000:])
