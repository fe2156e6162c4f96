"""cProfile of the python driver of a training step (host side): where the ~26 ms per step of enqueue time go.
    python tools/profile_train_host.py [steps]"""
import cProfile
import os
import pstats
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fcaf3d_b200.train import TrainStepModule, make_optimizer, train_step  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
from fcaf3d_b200 import me as ME  # noqa: E402
ME.set_compute_dtype(torch.bfloat16)
ctx = bench.Ctx()
det, scenes, _ = bench.make_model(ctx, "scannet", calibrate=False)
dev = ctx.dev
det.train()
module = TrainStepModule(det)
opt = make_optimizer(module)
pts = [torch.from_numpy(p).to(dev) for p, _, _ in scenes]
gtb = [torch.from_numpy(g).to(dev) for _, g, _ in scenes]
gtl = [torch.from_numpy(l).to(dev) for _, _, l in scenes]
for _ in range(3):
    train_step(module, opt, pts, gtb, gtl)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(steps):
    train_step(module, opt, pts, gtb, gtl)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(45)
st.sort_stats("cumulative").print_stats(40)
