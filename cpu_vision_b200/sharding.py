"""Shard-by-image multi-GPU harness.

RoIs and boxes belong to exactly one image and images are independent, so the
path shards with no data-path collective: rank r owns the contiguous image
block [r*B/W, (r+1)*B/W), its feature maps, its RoIs (batch index rebased to
the local block) and its per-image box sets.  The only exchange is optional
and sits AFTER the kernels: an all-gather that reassembles pooled features and
kept indices for a consumer that wants them on every rank.

The reference has no counterpart inside the library (its only collectives are
in references/detection/utils.py); these helpers are what its DDP recipes would
call around torchvision.ops.
"""
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def image_block(num_images: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[lo, hi) of the images rank owns; blocks differ by at most one image."""
    base, rem = divmod(num_images, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_rois(rois: torch.Tensor, num_images: int, rank: int, world_size: int):
    """Select this rank's RoIs from a global Tensor[K,5] and rebase their batch index.

    Returns (local_rois, global_row_index) -- the second lets the caller put
    local results back at the global positions.
    """
    lo, hi = image_block(num_images, rank, world_size)
    b = rois[:, 0]
    sel = torch.nonzero((b >= lo) & (b < hi), as_tuple=False).flatten()
    local = rois.index_select(0, sel).clone()
    local[:, 0] -= lo
    return local, sel


def all_gather_rows(local: torch.Tensor, counts: Sequence[int] = None, group=None) -> torch.Tensor:
    """Concatenate every rank's [k_r, ...] rows in rank order.

    Equal k_r -> one all_gather_into_tensor straight into the result (each rank
    can also have produced its rows in place: see gather_buffer()).  Unequal k_r
    -> pad to the maximum, gather, and drop the padding.
    """
    world = dist.get_world_size(group)
    if world == 1:
        return local
    local = local.contiguous()
    if counts is None:
        c = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
        allc = torch.empty(world, dtype=torch.int64, device=local.device)
        dist.all_gather_into_tensor(allc, c, group=group)
        counts = allc.tolist()
    kmax = max(counts)
    if all(c == kmax for c in counts):
        out = torch.empty((world * kmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local, group=group)
        return out
    padded = torch.zeros((kmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    out = torch.empty((world * kmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return torch.cat([out[r * kmax: r * kmax + counts[r]] for r in range(world)], dim=0)


def gather_buffer(k_local: int, tail_shape: Sequence[int], dtype, device, rank: int, world_size: int,
                  memory_format=torch.contiguous_format):
    """Allocate the full [W*k_local, ...] gather target and return (full, my_slice).

    The roi_align kernel writes straight into `my_slice`, so the all-gather
    needs no staging copy: all_gather_into_tensor(full, my_slice) is in place.
    """
    full = torch.empty((world_size * k_local,) + tuple(tail_shape), dtype=dtype, device=device,
                       memory_format=memory_format)
    return full, full[rank * k_local: (rank + 1) * k_local]


class _RawDeviceBytes:
    """CUDA array interface over a raw device range (zero-copy view for torch.as_tensor)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class PushGather:
    """The gather buffer of the pooled rows, written by every rank's forward kernel ON ALL RANKS (push all-gather).

    `full` is this rank's [W * k_local, C, PH, PW] channels_last buffer in peer-visible memory (a cudaMalloc of its
    own, exported through CUDA IPC and opened by every other rank of the node).  forward() runs
    b200det_msroi_align_forward_push: a pooled bin is stored to this rank's slice of `full` and, by the same
    kernel, to this rank's slice of every peer's buffer over NVLink -- the all-gather of north_star happens while
    the rows are produced, with no separate collective pass over the 6.4 GB (cfg4).  Two stream-ordered 4-byte
    all-reduces order the ranks: nobody overwrites a buffer a peer may still read, nobody reads a row before every
    kernel has finished.  Rows are ordered rank, RoI, exactly like all_gather_into_tensor of the local outputs
    (tests/test_multigpu_gpu.py: bit for bit)."""

    def __init__(self, k_local: int, channels: int, pooled_h: int, pooled_w: int, dtype, device, group=None):
        from . import capi

        self.capi, self.group, self.device = capi, group, torch.device(device)
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > 8:
            raise ValueError("PushGather: at most 8 ranks (one NVLink domain; the kernel pushes to 7 peers)")
        es = torch.empty(0, dtype=dtype).element_size()
        self.slice_bytes = k_local * channels * pooled_h * pooled_w * es
        nbytes = max(self.world * self.slice_bytes, 16)
        self.ptr = capi.peer_alloc(nbytes, self.device)
        handles = [None] * self.world
        dist.all_gather_object(handles, capi.peer_export(self.ptr, self.device), group=group)
        self.peer_base = {r: capi.peer_open(handles[r], self.device) for r in range(self.world) if r != self.rank}
        raw = torch.as_tensor(_RawDeviceBytes(self.ptr, nbytes), device=self.device)
        self.full = raw[: self.world * self.slice_bytes].view(dtype).view(
            self.world * k_local, pooled_h, pooled_w, channels).permute(0, 3, 1, 2)
        self.mine = self.full[self.rank * k_local: (self.rank + 1) * k_local]
        self.peer_ptrs = [b + self.rank * self.slice_bytes for _, b in sorted(self.peer_base.items())]
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.barrier()

    def barrier(self):
        """Stream-ordered: completes on a rank only after every rank's stream has reached it."""
        dist.all_reduce(self._flag, group=self.group)

    def forward(self, feats, rois, scales, pooled_h, pooled_w, sampling_ratio, aligned=False, levels=None, **kw):
        """Pools this rank's RoIs into every rank's buffer; returns (full, levels) once all ranks' rows are in."""
        self.barrier()
        _, levels = self.capi.msroi_align_forward_push(feats, rois, scales, pooled_h, pooled_w, sampling_ratio,
                                                       self.mine, self.peer_ptrs, aligned=aligned, levels=levels, **kw)
        self.barrier()
        return self.full, levels

    def close(self):
        if self.ptr is None:
            return
        torch.cuda.synchronize(self.device)
        self.barrier()
        torch.cuda.synchronize(self.device)
        for b in self.peer_base.values():
            self.capi.peer_close(b, self.device)
        self.full = self.mine = None
        dist.barrier(group=self.group)  # nobody frees while a peer still has the buffer mapped
        self.capi.peer_free(self.ptr, self.device)
        self.ptr = None


def all_gather_keep(keep: torch.Tensor, count: torch.Tensor, capacity: int, group=None) -> List[torch.Tensor]:
    """Gather variable-length kept-index lists as (count, padded int64[capacity]) per rank."""
    world = dist.get_world_size(group)
    pack = torch.zeros(capacity + 1, dtype=torch.int64, device=keep.device)
    pack[0] = count.reshape(())
    n = min(int(keep.shape[0]), capacity)
    pack[1: 1 + n] = keep[:n]
    if world == 1:
        return [pack[1: 1 + int(pack[0])]]
    out = torch.empty(world * (capacity + 1), dtype=torch.int64, device=keep.device)
    dist.all_gather_into_tensor(out, pack, group=group)
    out = out.view(world, capacity + 1)
    return [out[r, 1: 1 + int(out[r, 0])] for r in range(world)]
