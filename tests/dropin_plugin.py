"""pytest plugin for the drop-in runs (tests/test_dropin_gpu.py): imports the reference's torchvision from
baseline/_ref (first on PYTHONPATH) and THEN cpu_vision_b200, so that the reference's own tests run with its
schemas, Autograd, Meta and CPU kernels and with this repository's CUDA + Autocast kernels."""
import os

import torchvision

assert os.sep + os.path.join("baseline", "_ref", "torchvision") in torchvision.__file__, torchvision.__file__

import cpu_vision_b200  # noqa: E402,F401
from cpu_vision_b200 import _lib  # noqa: E402

assert (_lib.defined_schemas() & 3) == 0, "the reference's _C must have defined torchvision::roi_align / nms"
