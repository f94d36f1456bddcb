import os, sys
os.environ["B200DET_BWD_STRIP"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import oracle
from cpu_vision_b200 import capi
DEV = "cuda:0"
N, C, H, W, PH, PW, scale, sr, aligned = 1, 256, 50, 68, 7, 7, 0.0625, 2, True
rois = torch.tensor([[0.0, 174.68511962890625, 4.134002685546875, 400.21038818359375, 422.6258239746094]])
go = torch.ones(1, C, PH, PW)
want = oracle.roi_align_backward(go.double(), rois.double(), scale, PH, PW, N, C, H, W, sr, aligned)
gi = capi.roi_align_backward(go.to(DEV).contiguous(memory_format=torch.channels_last), rois.to(DEV), scale, PH, PW, N, C, H, W, sr, aligned, layout=capi.NHWC)
g = gi.cpu().double()
torch.set_printoptions(linewidth=250, precision=3, sci_mode=False)
print("col sums want", want[0, 0].sum(0)[8:28])
print("col sums got ", g[0, 0].sum(0)[8:28])
print("row sums want", want[0, 0].sum(1)[0:28])
print("row sums got ", g[0, 0].sum(1)[0:28])
print("ch diff", (g - want).abs().amax(dim=(0, 2, 3))[:8])
d = (g - want)[0, 0]
for r in range(0, 28):
    print(r, " ".join(f"{v:6.3f}" for v in d[r, 8:28].tolist()))
