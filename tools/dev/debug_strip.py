import os, sys
os.environ["B200DET_BWD_STRIP"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import oracle
from cpu_vision_b200 import capi
from test_roi_align_gpu import rand_rois, CASES
DEV = "cuda:0"
case = CASES[int(sys.argv[1]) if len(sys.argv) > 1 else 0]
N, C, H, W, K, PH, PW, scale, sr, aligned = case
g = torch.Generator().manual_seed(hash(case) % (2 ** 31))
x = torch.randn(N, C, H, W, generator=g)
rois = rand_rois(g, K, N, W / scale, H / scale, lo=1.0 / scale)
go = torch.randn(K, C, PH, PW, generator=g)
keep = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else list(range(K))
rois, go = rois[keep].contiguous(), go[keep].contiguous()
want = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PW, N, C, H, W, sr, aligned)
gi = capi.roi_align_backward(go.to(DEV).contiguous(memory_format=torch.channels_last), rois.to(DEV), scale, PH, PW, N, C, H, W, sr, aligned, layout=capi.NHWC)
d = (gi.cpu().double() - want).abs()
print("max err", d.max().item(), "of", want.abs().max().item())
bad = (d > 1e-3).nonzero()
print("bad elements", bad.shape[0])
pix = {}
for n, c, h, w in bad.tolist():
    pix.setdefault((n, h, w), 0); pix[(n, h, w)] += 1
for kx in sorted(pix)[:60]:
    print(kx, pix[kx])
# per RoI check
if len(keep) == K:
    for k in range(K):
        w1 = oracle.roi_align_backward(go[k:k+1].double(), rois[k:k+1].double(), scale, PH, PW, N, C, H, W, sr, aligned)
        g1 = capi.roi_align_backward(go[k:k+1].to(DEV).contiguous(memory_format=torch.channels_last), rois[k:k+1].to(DEV), scale, PH, PW, N, C, H, W, sr, aligned, layout=capi.NHWC)
        e = (g1.cpu().double() - w1).abs().max().item()
        if e > 1e-4:
            print("roi", k, rois[k].tolist(), "scaled", (rois[k, 1:] * scale).tolist(), "err", e)
