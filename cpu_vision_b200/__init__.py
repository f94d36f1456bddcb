"""cpu_vision_b200 -- B200-native (sm_100a) roi_align / nms hot path behind the
operator API of kaisun24/CPU-Vision (a torchvision 0.20 fork).

Public surface (same names and signatures as the reference's torchvision.ops):
    roi_align, RoIAlign, MultiScaleRoIAlign, nms, batched_nms, roi_pool, RoIPool, ps_roi_pool, PSRoIPool, ps_roi_align, PSRoIAlign
    cpu_vision_b200.detection    filter_proposals / postprocess_detections batched over images
    cpu_vision_b200.dropin       enable() / disable(): the reference's detection classes on the mirrors, at run time
Lower level:
    cpu_vision_b200.capi         ctypes binding of the C-ABI (include/b200_detection_ops.h)
    cpu_vision_b200.sharding     shard-by-image multi-GPU harness (NCCL all-gather)
Importing the package loads libb200det.so and registers the CUDA kernels for
torch.ops.torchvision.{roi_align,_roi_align_backward,nms}; it raises ImportError
when the native libraries are not built.  There is no CPU fallback.
"""
from . import _lib

_lib.load()

from .boxes import batched_nms, nms  # noqa: E402
from .poolers import LevelMapper, MultiScaleRoIAlign  # noqa: E402
from .roi_align import RoIAlign, roi_align  # noqa: E402
from .roi_pool import PSRoIAlign, PSRoIPool, RoIPool, ps_roi_align, ps_roi_pool, roi_pool  # noqa: E402

__all__ = ["roi_align", "RoIAlign", "MultiScaleRoIAlign", "LevelMapper", "nms", "batched_nms", "roi_pool", "RoIPool", "ps_roi_pool", "PSRoIPool", "ps_roi_align", "PSRoIAlign"]
__version__ = "0.1.0"
