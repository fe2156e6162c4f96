"""Diagnostic: two forward/backward passes of the training step on the same batch - which parameter gradients differ, by how much
(relative to the largest entry of that gradient)?   python tools/train_determinism.py [scannet|sunrgbd]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fcaf3d_b200 import me as ME  # noqa: E402
from fcaf3d_b200.detector import build_fcaf3d  # noqa: E402
from fcaf3d_b200.scenes import make_scene  # noqa: E402
from fcaf3d_b200.testing import deterministic_init  # noqa: E402

ds = sys.argv[1] if len(sys.argv) > 1 else "scannet"
dev = torch.device("cuda", 0)
ME.set_compute_dtype(torch.bfloat16)
ncls = {"scannet": 18, "sunrgbd": 10}[ds]
det = deterministic_init(build_fcaf3d(ds), 0, cls_bias=0.0).to(dev).train()
sc = [make_scene("S0" if ds == "scannet" else "S3", 100 * i, n_classes=ncls) for i in range(8)]
pts = [torch.from_numpy(p).to(dev) for p, _, _ in sc]
gtb = [torch.from_numpy(g).to(dev) for _, g, _ in sc]
gtl = [torch.from_numpy(l).to(dev) for _, _, l in sc]
grads = []
for it in range(3):
    det.zero_grad(set_to_none=True)
    losses = det.forward_train(pts, gtb, gtl)
    sum(losses.values()).backward()
    grads.append({n: p.grad.detach().clone() for n, p in det.named_parameters()})
    print(it, {k: round(float(v), 6) for k, v in losses.items()})
worst = []
for n in grads[0]:
    a, b = grads[0][n], grads[1][n]
    d = ((a - b).abs().max() / a.abs().max().clamp(min=1e-30)).item()
    worst.append((d, n, float(a.abs().max())))
worst.sort(reverse=True)
for d, n, m in worst[:12]:
    print("%.3e  %-50s max|g| %.3e" % (d, n, m))
