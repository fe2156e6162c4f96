"""Throughput of the device-side data path (fcaf3d_b200/data.py) next to the numpy restatement of the reference's
transforms (oracle/data_cpu.py) on the host: 8 raw scenes of 150 000 points -> sample 100 000, flip, rotate/scale/translate.
    python tools/bench_data.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fcaf3d_b200.data import DeviceAugment, ScenePrefetcher  # noqa: E402
from oracle import data_cpu as D  # noqa: E402

dev = torch.device("cuda", 0)
rng = np.random.default_rng(0)
raw = [rng.uniform(-4, 4, (150000, 6)).astype(np.float32) for _ in range(8)]
gt = [np.concatenate([rng.uniform(-3, 3, (15, 3)), rng.uniform(.3, 2, (15, 3)), np.zeros((15, 1))], 1).astype(np.float32)
      for _ in range(8)]
# host pipeline (what 4 dataloader workers per GPU run in the reference), one thread
t0 = time.perf_counter()
reps = 3
for _ in range(reps):
    for p, b in zip(raw, gt):
        pr = D.draw_params(rng, len(p), 100000)
        q, c = D.random_flip(p[pr["choices"]], b, pr["flip_h"], pr["flip_v"], False)
        D.global_rot_scale_trans(q, c, pr["angle"], pr["scale"], pr["trans"], False)
cpu_ms = (time.perf_counter() - t0) / reps * 1e3
# device pipeline, scenes resident
aug = DeviceAugment(num_points=100000, with_yaw=False, seed=0)
dp = [torch.from_numpy(p).to(dev) for p in raw]
db = [torch.from_numpy(b).to(dev) for b in gt]
for _ in range(3):
    aug(dp, db)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(20):
    aug(dp, db)
e.record()
torch.cuda.synchronize()
gpu_ms = s.elapsed_time(e) / 20
# end to end: pinned host -> device through the prefetcher, then augment
pinned = [[torch.from_numpy(p).pin_memory() for p in raw] for _ in range(12)]
torch.cuda.synchronize()
t0 = time.perf_counter()
n = 0
for batch in ScenePrefetcher(pinned, dev):
    aug(batch, db)
    n += 1
torch.cuda.synchronize()
e2e_ms = (time.perf_counter() - t0) / n * 1e3
print("8 scenes x 150k -> 100k points: numpy host pipeline %.1f ms (1 thread) | device %.2f ms | pinned->device+augment %.2f ms"
      % (cpu_ms, gpu_ms, e2e_ms))
print("scenes/s: host %.0f, device %.0f, end to end %.0f" % (8e3 / cpu_ms, 8e3 / gpu_ms, 8e3 / e2e_ms))
