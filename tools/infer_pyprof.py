"""cProfile of the host side of the fused inference step (diagnostic): python tools/infer_pyprof.py"""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fcaf3d_b200 import me as ME  # noqa: E402
from fcaf3d_b200.detector import build_fcaf3d  # noqa: E402
from fcaf3d_b200.scenes import make_scene  # noqa: E402
from fcaf3d_b200.testing import deterministic_init  # noqa: E402

dev = torch.device("cuda", 0)
ME.set_compute_dtype(torch.bfloat16)
det = deterministic_init(build_fcaf3d("scannet"), 0, cls_bias=0.0).to(dev)
pts = [torch.from_numpy(make_scene("S0", 100 * i, n_classes=18)[0]).to(dev) for i in range(8)]
with torch.no_grad():
    det.train()
    for m in det.modules():
        if isinstance(m, torch.nn.BatchNorm1d):
            m.momentum = 1.0
    det.forward_features_fused(pts)
    det.eval()
    for _ in range(3):
        det.simple_test_fused(pts, to_host=False)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(20):
        det.simple_test_fused(pts, to_host=False)
    torch.cuda.synchronize()
    pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(45)
