"""Where the host time of one training step goes (torch.profiler; diagnostic, not a bench).
    python tools/train_hosttime.py > gpurun_out/train_prof.txt"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fcaf3d_b200 import me as ME  # noqa: E402
from fcaf3d_b200.detector import build_fcaf3d  # noqa: E402
from fcaf3d_b200.scenes import make_scene  # noqa: E402
from fcaf3d_b200.testing import deterministic_init  # noqa: E402
from fcaf3d_b200.train import TrainStepModule, make_optimizer, train_step  # noqa: E402

dev = torch.device("cuda", 0)
ME.set_compute_dtype(torch.bfloat16)
det = deterministic_init(build_fcaf3d("scannet"), 0, cls_bias=0.0).to(dev)
det.train()
module = TrainStepModule(det)
opt = make_optimizer(module)
scenes = [make_scene("S0", 100 * i, n_classes=18) for i in range(8)]
pts = [torch.from_numpy(p).to(dev) for p, _, _ in scenes]
gtb = [torch.from_numpy(g).to(dev) for _, g, _ in scenes]
gtl = [torch.from_numpy(l).to(dev) for _, _, l in scenes]
for _ in range(3):
    train_step(module, opt, pts, gtb, gtl)
torch.cuda.synchronize()
t0 = time.perf_counter()
train_step(module, opt, pts, gtb, gtl)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host enqueue %.1f ms, + drain %.1f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3))
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    train_step(module, opt, pts, gtb, gtl)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=40, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=30, max_name_column_width=60))
