"""Generates tests/golden/roi_pool.pt and ps_roi_pool.pt from the REFERENCE's own torchvision.ops.roi_pool /
ps_roi_pool (compiled CPU kernels, autograd included), built in /tmp by make_golden.build_reference.

    python tests/golden/make_golden_roi_pool.py
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden  # noqa: E402

GEN = r'''
import sys, torch
sys.path.insert(0, "%(work)s")
import torchvision
assert torchvision.__file__.startswith("%(work)s"), torchvision.__file__
from torchvision import ops
g = torch.Generator().manual_seed(20261106)
x = torch.randn(2, 6, 23, 31, generator=g)
k = 40
cx = torch.rand(k, generator=g) * 124; cy = torch.rand(k, generator=g) * 92
w = torch.rand(k, generator=g) * 80 + 1; h = torch.rand(k, generator=g) * 60 + 1
rois = torch.stack([torch.randint(0, 2, (k,), generator=g).float(), cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
rois[0, 1:] = torch.tensor([50.0, 60.0, 40.0, 30.0])      # malformed
rois[1, 1:] = torch.tensor([-40.0, -40.0, -8.0, -8.0])    # outside: empty bins
rois[2, 1:] = torch.tensor([100.0, 80.0, 300.0, 300.0])   # beyond the far border
rois[3, 1:] = torch.tensor([30.0, 30.0, 30.0, 30.0])      # one pixel
cases = []
for (ph, pw) in ((7, 7), (3, 5), (1, 1)):
    for scale in (0.25, 1.0 / 3.0):
        xi = x.clone().requires_grad_(True)
        y = ops.roi_pool(xi, rois, (ph, pw), scale)
        gy = torch.randn(y.shape, generator=g)
        y.backward(gy)
        cases.append(dict(ph=ph, pw=pw, scale=scale, out=y.detach().clone(), grad_out=gy, grad_in=xi.grad.clone()))
torch.save(dict(x=x, rois=rois, cases=cases), "%(out)s/roi_pool.pt")
print("roi_pool golden written", len(cases))
# ---- ps_roi_pool: channels must be a multiple of PH * PW
pcases = []
for (ph, pw, cmul) in ((3, 3, 2), (2, 4, 3), (1, 1, 5)):
    xp = torch.randn(2, cmul * ph * pw, 23, 31, generator=g)
    for scale in (0.25, 1.0 / 3.0):
        xi = xp.clone().requires_grad_(True)
        y = ops.ps_roi_pool(xi, rois, (ph, pw), scale)
        gy = torch.randn(y.shape, generator=g)
        y.backward(gy)
        pcases.append(dict(ph=ph, pw=pw, scale=scale, x=xp, out=y.detach().clone(), grad_out=gy, grad_in=xi.grad.clone()))
torch.save(dict(rois=rois, cases=pcases), "%(out)s/ps_roi_pool.pt")
print("ps_roi_pool golden written", len(pcases))
acases = []
for (ph, pw, cmul) in ((3, 3, 2), (2, 4, 3)):
    xp = torch.randn(2, cmul * ph * pw, 23, 31, generator=g)
    for (scale, sr) in ((0.25, 2), (1.0 / 3.0, -1), (0.25, 3)):
        xi = xp.clone().requires_grad_(True)
        y = ops.ps_roi_align(xi, rois, (ph, pw), scale, sr)
        gy = torch.randn(y.shape, generator=g)
        y.backward(gy)
        acases.append(dict(ph=ph, pw=pw, scale=scale, sr=sr, x=xp, out=y.detach().clone(), grad_out=gy,
                           grad_in=xi.grad.clone()))
torch.save(dict(rois=rois, cases=acases), "%(out)s/ps_roi_align.pt")
print("ps_roi_align golden written", len(acases))
'''


def main():
    make_golden.build_reference()
    subprocess.check_call([sys.executable, "-c", GEN % dict(work=make_golden.WORK, out=HERE)])


if __name__ == "__main__":
    main()
