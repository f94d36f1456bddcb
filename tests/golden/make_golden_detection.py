"""Generates tests/golden/detection.pt from the REFERENCE's own RegionProposalNetwork.filter_proposals
(models/detection/rpn.py:242-297) and RoIHeads.postprocess_detections (roi_heads.py:668-725), run on CPU
tensors through the reference's compiled CPU ops (built in /tmp by make_golden.build_reference; nothing
from the reference is written into this repository).

    python tests/golden/make_golden_detection.py
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
from torchvision.models.detection.rpn import RegionProposalNetwork, RPNHead
from torchvision.models.detection.anchor_utils import AnchorGenerator
from torchvision.models.detection.roi_heads import RoIHeads
g = torch.Generator().manual_seed(20261105)
out = {}

def boxes(n, w, h, lo, hi):
    cx = torch.rand(n, generator=g) * w; cy = torch.rand(n, generator=g) * h
    bw = torch.rand(n, generator=g) * (hi - lo) + lo; bh = torch.rand(n, generator=g) * (hi - lo) + lo
    # some boxes reach outside the image (clipped), some collapse (removed as small)
    return torch.stack([cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2], 1)

# ---- RPN filter_proposals: 3 images, 5 levels ------------------------------------------------------
image_shapes = [(320, 416), (300, 400), (256, 416)]
napl = [4800, 1200, 300, 75, 27]
A = sum(napl)
props = torch.stack([boxes(A, 416.0, 320.0, 0.0, 150.0) for _ in image_shapes])
props[:, ::97] = props[:, ::97].round()            # exact duplicates / zero-size boxes after clipping
props[:, 5::211, 2:] = props[:, 5::211, :2]        # zero-size
obj = torch.randn(len(image_shapes) * A, 1, generator=g) * 2
cases = []
for (pre, post, thr, sthr, msz) in ((300, 200, 0.7, 0.0, 1e-3), (1000, 5000, 0.5, 0.3, 4.0), (50, 20, 0.7, 0.0, 1e-3)):
    rpn = RegionProposalNetwork(AnchorGenerator(), RPNHead(8, 3), 0.7, 0.3, 256, 0.5,
                                dict(training=pre, testing=pre), dict(training=post, testing=post), thr,
                                score_thresh=sthr)
    rpn.min_size = msz
    rpn.eval()
    b, s = rpn.filter_proposals(props.clone(), obj.clone(), image_shapes, napl)
    cases.append(dict(pre=pre, post=post, nms_thresh=thr, score_thresh=sthr, min_size=msz, boxes=b, scores=s))
out["rpn"] = dict(proposals=props, objectness=obj, image_shapes=image_shapes, num_anchors_per_level=napl, cases=cases)

# ---- RoIHeads.postprocess_detections: 3 images, 11 classes ----------------------------------------------
per = [120, 77, 150]
ncls = 11
proposals = [boxes(n, 416.0, 320.0, 8.0, 200.0).clamp(min=0) for n in per]
R = sum(per)
logits = torch.randn(R, ncls, generator=g) * 3
reg = torch.randn(R, ncls * 4, generator=g)
reg[::13] *= 40.0                                  # exercises the bbox_xform_clip clamp
dcases = []
for (sthr, nthr, dpi) in ((0.05, 0.5, 100), (0.2, 0.3, 10), (0.0, 0.7, 1000)):
    rh = RoIHeads(None, None, None, 0.5, 0.5, 512, 0.25, None, sthr, nthr, dpi)
    b, s, l = rh.postprocess_detections(logits.clone(), reg.clone(), [p.clone() for p in proposals], image_shapes)
    dcases.append(dict(score_thresh=sthr, nms_thresh=nthr, detections_per_img=dpi, boxes=b, scores=s, labels=l))
out["roi_heads"] = dict(class_logits=logits, box_regression=reg, proposals=proposals, image_shapes=image_shapes,
                        cases=dcases)
torch.save(out, "%(out)s/detection.pt")
print("detection golden written:", [[t.shape[0] for t in c["boxes"]] for c in cases],
      [[t.shape[0] for t in c["boxes"]] for c in dcases])
'''


def main():
    make_golden.build_reference()
    code = GEN % dict(work=make_golden.WORK, out=HERE)
    subprocess.check_call([sys.executable, "-c", code])


if __name__ == "__main__":
    main()
