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
PIPE = os.environ.get("PIPELINED", "0") == "1"
if PIPE:   # steady state of the software-pipelined trainer: profile the 2nd and 3rd of four chained steps
    prepared = det.prepare(pts)
    _, _, prepared = train_step(module, opt, prepared, gtb, gtl, next_points=pts)
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    if PIPE:
        _, _, prepared = train_step(module, opt, prepared, gtb, gtl, next_points=pts)
        _, _, prepared = train_step(module, opt, prepared, gtb, gtl, next_points=pts)
    else:
        train_step(module, opt, pts, gtb, gtl)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=40, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=30, max_name_column_width=60))
# GPU idle gaps: sort kernel events by start, list the largest gaps and what ran before / after
evs = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range is not None],
             key=lambda e: e.time_range.start)
gaps = []
for a, b in zip(evs[:-1], evs[1:]):
    g = b.time_range.start - a.time_range.end
    if g > 0:
        gaps.append((g, a.name[:60], b.name[:60], a.time_range.end - evs[0].time_range.start))
tot = sum(g[0] for g in gaps)
span = evs[-1].time_range.end - evs[0].time_range.start
busy = sum(e.time_range.end - e.time_range.start for e in evs)
print("GPU span %.2f ms, busy %.2f ms, idle %.2f ms in %d gaps" % (span / 1e3, busy / 1e3, tot / 1e3, len(gaps)))
for g, a, b, t in sorted(gaps, reverse=True)[:40]:
    print("  gap %7.1f us at %6.2f ms  after %-60s before %s" % (g, t / 1e3, a, b))
# idle time per 2 ms window of the step
import collections
win = collections.Counter()
for g, a, b, t in gaps:
    win[int(t // 2000)] += g
print("idle us per 2 ms window:", [int(win[i]) for i in range(int(span // 2000) + 1)])
