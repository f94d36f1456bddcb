"""The exchange north_star names, on real GPUs over NCCL: images sharded by rank, roi_align forward written into
each rank's slice of the gather buffer, NCCL all-gather of the pooled features and of the kept indices -- the
result must equal the one-GPU result BIT FOR BIT (the reference has no such test: its only collectives live in
references/detection/utils.py).  The fused variant (b200det_msroi_align_forward_push through sharding.PushGather:
the forward kernel stores every pooled row into all ranks' buffers over NVLink peer memory) must leave the same
bytes on every rank.  Needs >= 2 GPUs; skipped otherwise (gloo twin: tests/test_host_logic.py)."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(100, 136), (50, 68), (25, 34)]
SCALES = [0.125, 0.0625, 0.03125]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _inputs(images, per_img):
    g = torch.Generator().manual_seed(77)
    feats = [torch.randn(images, 256, h, w, generator=g) for h, w in SHAPES]
    ctr = torch.rand(images * per_img, 2, generator=g) * torch.tensor([1088.0, 800.0])
    wh = torch.exp(torch.rand(images * per_img, 2, generator=g) * 3.0 + 3.0)
    b = torch.arange(images).repeat_interleave(per_img).float()
    rois = torch.cat([b[:, None], (ctr - wh / 2).clamp(min=0), (ctr + wh / 2)], 1)
    boxes, scores, idxs = [], [], []
    for i in range(images):
        xy = torch.rand(1500, 2, generator=g) * 500
        boxes.append(torch.cat([xy, xy + torch.rand(1500, 2, generator=g) * 100 + 10], 1))
        scores.append(torch.randperm(1500, generator=g).float() / 1500)
        idxs.append(torch.randint(0, 20, (1500,), generator=g))
    return feats, rois, boxes, scores, idxs


def _worker(rank, world, port, images, per_img, ret):
    import torch.distributed as dist

    from cpu_vision_b200 import capi, sharding

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        feats, rois, boxes, scores, idxs = _inputs(images, per_img)
        lo, hi = sharding.image_block(images, rank, world)
        my_feats = [f[lo:hi].to(dev).contiguous(memory_format=torch.channels_last) for f in feats]
        my_rois, _ = sharding.shard_rois(rois.to(dev), images, rank, world)
        k_local = my_rois.shape[0]
        full, mine = sharding.gather_buffer(k_local, (256, 7, 7), torch.float32, dev, rank, world, torch.channels_last)
        capi.msroi_align_forward(my_feats, my_rois, SCALES, 7, 7, 2, False, k_min=3, k_max=5, out=mine)
        dist.all_gather_into_tensor(full, mine)
        # the same exchange fused into the forward: every rank's kernel stores its rows into all ranks' buffers
        pg = sharding.PushGather(k_local, 256, 7, 7, torch.float32, dev)
        pushed, _ = pg.forward(my_feats, my_rois, SCALES, 7, 7, 2, False, k_min=3, k_max=5)
        torch.cuda.synchronize()
        push_ok = torch.equal(pushed, full)
        pushed2, _ = pg.forward(my_feats, my_rois, SCALES, 7, 7, 2, False, k_min=3, k_max=5)  # buffer reuse
        torch.cuda.synchronize()
        push_ok = push_ok and torch.equal(pushed2, full)
        flags = [None] * world
        dist.all_gather_object(flags, bool(push_ok))
        pg.close()
        kept = []
        for i in range(lo, hi):  # kept indices of this rank's images, gathered image by image
            keep = capi.batched_nms(boxes[i].to(dev), scores[i].to(dev), idxs[i].to(dev), 0.5)
            kept.append(sharding.all_gather_keep(keep, torch.tensor([keep.numel()], device=dev), 1500))
        torch.cuda.synchronize()
        if rank == 0:
            # the one-GPU result of the whole batch
            all_feats = [f.to(dev).contiguous(memory_format=torch.channels_last) for f in feats]
            want, _ = capi.msroi_align_forward(all_feats, rois.to(dev), SCALES, 7, 7, 2, False, k_min=3, k_max=5)
            ok = torch.equal(full, want)
            per_rank = images // world
            for j in range(per_rank):
                for r in range(world):
                    i = r * per_rank + j
                    ref = capi.batched_nms(boxes[i].to(dev), scores[i].to(dev), idxs[i].to(dev), 0.5)
                    ok = ok and torch.equal(kept[j][r], ref)
            ret["ok"] = bool(ok)
            ret["push_ok"] = all(flags)
            ret["rows"] = int(full.shape[0])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gathered_pooled_features_and_kept_indices_equal_the_one_gpu_result(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp

    images, per_img = 8, 300
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, _free_port(), images, per_img, ret), nprocs=world, join=True)
        assert ret.get("ok") is True, dict(ret)
        assert ret.get("push_ok") is True, dict(ret)   # the fused forward + push gather, on every rank
        assert ret["rows"] == images * per_img
