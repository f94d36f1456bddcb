"""Generates tests/golden/*.pt from the REFERENCE ITSELF (run in the build container only).

    python tests/golden/make_golden.py

1. Builds the reference's CPU-only `torchvision._C` in /tmp from the sources
   where they lie under /root/reference (the list setup.py:155-162 uses:
   csrc/*.cpp, csrc/ops/*.cpp, csrc/ops/{autocast,autograd,cpu,quantized/cpu}/*.cpp),
   one g++ call per file, and copies the reference's Python package next to it.
   Nothing from the reference is written into this repository.
2. Imports THAT torchvision in a subprocess and records inputs + outputs of
   roi_align (fwd/bwd), nms, batched_nms (both strategies), LevelMapper and
   MultiScaleRoIAlign on small seeded cases, plus the known-answer vectors
   K1-K7 of SURVEY.md section 8(a).
The fixtures are committed; /root/reference is not needed to run the tests.
"""
import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

REF = "/root/reference"
WORK = "/tmp/tvref_golden"
HERE = os.path.dirname(os.path.abspath(__file__))


def build_reference():
    import torch
    from torch.utils import cpp_extension as ce

    pkg = os.path.join(WORK, "torchvision")
    so = os.path.join(pkg, "_C.so")
    if os.path.exists(so):
        return
    if os.path.exists(WORK):
        shutil.rmtree(WORK)
    os.makedirs(os.path.join(WORK, "obj"))
    shutil.copytree(os.path.join(REF, "torchvision"), pkg,
                    ignore=shutil.ignore_patterns("csrc", "*.so", "__pycache__"))
    with open(os.path.join(pkg, "version.py"), "w") as f:
        f.write("__version__ = '0.20.0a0+ref'\ngit_version = 'reference'\n"
                "from torchvision.extension import _check_cuda_version\n"
                "if _check_cuda_version() > 0:\n    cuda = _check_cuda_version()\n")
    csrc = os.path.join(REF, "torchvision", "csrc")
    srcs = glob.glob(os.path.join(csrc, "*.cpp")) + glob.glob(os.path.join(csrc, "ops", "*.cpp"))
    for sub in ("autocast", "autograd", "cpu", "quantized/cpu"):
        srcs += glob.glob(os.path.join(csrc, "ops", sub, "*.cpp"))
    inc = ["-I" + p for p in ce.include_paths()] + ["-I" + csrc]
    import sysconfig

    inc.append("-I" + sysconfig.get_paths()["include"])
    abi = "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)

    def cc(src):
        obj = os.path.join(WORK, "obj", os.path.relpath(src, csrc).replace("/", "__") + ".o")
        subprocess.check_call(["g++", "-O3", "-std=c++17", "-fPIC", abi, "-DTORCH_EXTENSION_NAME=_C"] + inc +
                              ["-c", src, "-o", obj])
        return obj

    with cf.ThreadPoolExecutor(8) as ex:
        objs = list(ex.map(cc, srcs))
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    subprocess.check_call(["g++", "-shared", "-o", so] + objs +
                          ["-L" + torch_lib, "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python",
                           "-Wl,-rpath," + torch_lib])


GEN = r'''
import sys, torch
sys.path.insert(0, "%(work)s")
import torchvision
assert torchvision.__file__.startswith("%(work)s"), torchvision.__file__
from torchvision import ops
from torchvision.ops import boxes as box_ops
from torchvision.ops.poolers import MultiScaleRoIAlign, LevelMapper
from collections import OrderedDict
out = "%(out)s"
g = torch.Generator().manual_seed(20261104)

def rnd_rois(k, n, w, h, lo=2.0, hi=None):
    hi = hi or min(w, h)
    cx = torch.rand(k, generator=g) * w; cy = torch.rand(k, generator=g) * h
    bw = torch.exp(torch.rand(k, generator=g) * (torch.log(torch.tensor(hi)) - torch.log(torch.tensor(lo))) + torch.log(torch.tensor(lo)))
    bh = torch.exp(torch.rand(k, generator=g) * (torch.log(torch.tensor(hi)) - torch.log(torch.tensor(lo))) + torch.log(torch.tensor(lo)))
    x1 = (cx - bw / 2).clamp(0, w - 1); y1 = (cy - bh / 2).clamp(0, h - 1)
    x2 = (cx + bw / 2).clamp(0, w); y2 = (cy + bh / 2).clamp(0, h)
    b = torch.randint(0, n, (k,), generator=g).float()
    return torch.stack([b, x1, y1, x2, y2], 1)

# ---- roi_align forward / backward --------------------------------------------------
cases = []
x = torch.randn(2, 12, 25, 31, generator=g)
rois = rnd_rois(24, 2, 124.0, 100.0)
rois[0, 1:] = torch.tensor([50.0, 60.0, 40.0, 30.0])      # malformed: x2<x1, y2<y1
rois[1, 1:] = torch.tensor([-20.0, -20.0, 10.0, 10.0])    # partly outside
rois[2, 1:] = torch.tensor([110.0, 90.0, 200.0, 180.0])   # beyond the far border
rois[3, 1:] = torch.tensor([30.0, 30.0, 30.0, 30.0])      # zero size
for (ph, pw) in ((7, 7), (3, 5)):
    for sr in (2, -1, 3):
        for aligned in (True, False):
            xi = x.clone().requires_grad_(True)
            y = ops.roi_align(xi, rois, (ph, pw), 0.25, sr, aligned)
            gy = torch.randn(y.shape, generator=g)
            y.backward(gy)
            cases.append(dict(ph=ph, pw=pw, scale=0.25, sr=sr, aligned=aligned, out=y.detach().clone(),
                              grad_out=gy, grad_in=xi.grad.clone()))
torch.save(dict(x=x, rois=rois, cases=cases), out + "/roi_align.pt")

# ---- known answers K1-K5 (fixtures of test/test_ops.py, test/test_onnx.py) -------
kat = {}
b = torch.tensor([[285.3538, 185.5758, 1193.5110, 851.4551], [285.1472, 188.7374, 1192.4984, 851.0669],
                  [279.2440, 197.9812, 1189.4746, 849.2019]])
s = torch.tensor([0.6370, 0.7569, 0.3966])
kat["K1_boxes"], kat["K1_scores"], kat["K1_keep"] = b, s, ops.nms(b, s, 0.2)
x8 = torch.linspace(1, 64, 64).reshape(1, 1, 8, 8)
r8 = torch.tensor([[0, 0, 0, 3, 3.0]])
kat["K2_x"], kat["K2_roi"] = x8, r8
kat["K2_unaligned_sr2"] = ops.roi_align(x8, r8, 2, 1.0, 2, False)
kat["K2_unaligned_srm1"] = ops.roi_align(x8, r8, 2, 1.0, -1, False)
kat["K2_aligned_sr2"] = ops.roi_align(x8, r8, 2, 1.0, 2, True)
x10 = torch.arange(100.0).reshape(1, 1, 10, 10)
r10 = torch.tensor([[0, 0.2, 0.3, 4.5, 3.5]])
kat["K3_x"], kat["K3_roi"] = x10, r10
kat["K3_a"] = ops.roi_align(x10, r10, 2, 2.5, 0, True)
kat["K3_b"] = ops.roi_align(x10, r10, 5, 0.5, 3, True)
r4 = torch.tensor([[0, 2, 0.3, 1.5, 1.5]])
kat["K4_roi"] = r4
kat["K4_aligned"] = ops.roi_align(x10, r4, 2, 1.0, 1, True)
kat["K4_unaligned"] = ops.roi_align(x10, r4, 2, 1.0, 1, False)
xi = x8.clone().requires_grad_(True)
ops.roi_align(xi, r8, 2, 1.0, 2, True).sum().backward()
kat["K5_grad"] = xi.grad.clone()
# K6: LevelMapper(2,5) on squares of side s
sides = torch.tensor([10.0, 111, 112, 113, 223, 224, 225, 447, 448, 800, 2000])
sq = torch.stack([torch.zeros_like(sides), torch.zeros_like(sides), sides, sides], 1)
kat["K6_boxes"], kat["K6_levels"] = sq, LevelMapper(2, 5)([sq])
torch.save(kat, out + "/known_answers.pt")

# ---- nms / batched_nms -----------------------------------------------------------
def boxes_scores(n, span, size_lo, size_hi):
    xy = torch.rand(n, 2, generator=g) * span
    wh = torch.rand(n, 2, generator=g) * (size_hi - size_lo) + size_lo
    sc = torch.randperm(n, generator=g).float() / n
    return torch.cat([xy, xy + wh], 1), sc
nm = {}
b, s = boxes_scores(600, 200.0, 10.0, 60.0)
nm["boxes"], nm["scores"] = b, s
for thr in (0.2, 0.5, 0.7):
    nm["keep_%%s" %% thr] = ops.nms(b, s, thr)
st = s.clone(); st[::3] = 0.5            # heavy ties: stable order must match
nm["scores_ties"], nm["keep_ties"] = st, ops.nms(b, st, 0.5)
idx = torch.randint(0, 7, (600,), generator=g)
nm["idxs"] = idx
nm["bn_trick"] = box_ops._batched_nms_coordinate_trick(b, s, idx, 0.5)
nm["bn_vanilla"] = box_ops._batched_nms_vanilla(b, s, idx, 0.5)
# large offsets where the coordinate trick rounds box coordinates (80 classes)
b2, s2 = boxes_scores(800, 1000.0, 10.0, 210.0)
idx2 = torch.randint(0, 80, (800,), generator=g)
nm["boxes2"], nm["scores2"], nm["idxs2"] = b2, s2, idx2
nm["bn2_trick"] = box_ops._batched_nms_coordinate_trick(b2, s2, idx2, 0.5)
nm["bn2_vanilla"] = box_ops._batched_nms_vanilla(b2, s2, idx2, 0.5)
torch.save(nm, out + "/nms.pt")

# ---- MultiScaleRoIAlign ------------------------------------------------------------
ms = {}
img = (480, 640)
feats = OrderedDict((str(i), torch.randn(2, 3, img[0] // (4 << i), img[1] // (4 << i), generator=g)) for i in range(4))
boxes = []
for i in range(2):
    r = rnd_rois(40, 1, float(img[1]), float(img[0]), lo=8.0, hi=900.0)[:, 1:]
    r[:, 2:] = torch.maximum(r[:, 2:], r[:, :2] + 1.0)
    big = torch.tensor([[5.0, 8.0, 630.0, 475.0], [100.0, 10.0, 560.0, 470.0], [0.0, 0.0, 448.0, 448.0],
                        [200.0, 100.0, 424.0, 324.0], [10.0, 10.0, 122.0, 122.0], [300.0, 200.0, 530.0, 420.0]])
    r = torch.cat([r, big + float(i)], 0)[torch.randperm(46, generator=g)]
    boxes.append(r)
for osz in (7, 14):
    m = MultiScaleRoIAlign(["0", "1", "2", "3"], osz, 2)
    fi = OrderedDict((k, v.clone().requires_grad_(True)) for k, v in feats.items())
    y = m(fi, boxes, [img, img])
    gy = torch.randn(y.shape, generator=g)
    y.backward(gy)
    ms["out_%%d" %% osz], ms["grad_out_%%d" %% osz] = y.detach().clone(), gy
    ms["grad_in_%%d" %% osz] = [fi[k].grad.clone() for k in fi]
    ms["scales"], ms["k_min"], ms["k_max"] = m.scales, m.map_levels.k_min, m.map_levels.k_max
ms["levels"] = m.map_levels(boxes)
ms["feats"], ms["boxes"], ms["image_shapes"] = list(feats.values()), boxes, [img, img]
# K7: scales inferred for the 800x1088 pyramid
m7 = MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2)
f7 = OrderedDict((str(i), torch.zeros(1, 1, 800 // (4 << i), 1088 // (4 << i))) for i in range(4))
m7(f7, [torch.tensor([[0.0, 0, 10, 10]])], [(800, 1088)])
ms["K7_scales"], ms["K7_kmin"], ms["K7_kmax"] = m7.scales, m7.map_levels.k_min, m7.map_levels.k_max
torch.save(ms, out + "/msroi_align.pt")
print("golden vectors written to", out)
'''


def main():
    build_reference()
    code = GEN % dict(work=WORK, out=HERE)
    subprocess.check_call([sys.executable, "-c", code], cwd="/tmp")
    for f in sorted(glob.glob(os.path.join(HERE, "*.pt"))):
        print(os.path.basename(f), os.path.getsize(f), "bytes")


if __name__ == "__main__":
    main()
