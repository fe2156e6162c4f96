"""Where the host time of one fused inference step goes (torch.profiler; diagnostic, not a bench).
    python tools/infer_hosttime.py > gpurun_out/infer_prof.txt"""
import os
import sys
import time

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
    t0 = time.perf_counter()
    det.simple_test_fused(pts, to_host=False)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("host enqueue %.2f ms, + drain %.2f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3))
    from torch.profiler import ProfilerActivity, profile  # noqa: E402
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        det.simple_test_fused(pts, to_host=False)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=25, max_name_column_width=60))
    # GPU idle gaps: sort kernel events by start, list the largest gaps and what ran before / after
    evs = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range is not None],
                 key=lambda e: e.time_range.start)
    gaps = []
    for a, b in zip(evs[:-1], evs[1:]):
        g = b.time_range.start - a.time_range.end
        if g > 0:
            gaps.append((g, a.name[:50], b.name[:50]))
    tot = sum(g for g, _, _ in gaps)
    span = evs[-1].time_range.end - evs[0].time_range.start
    busy = sum(e.time_range.end - e.time_range.start for e in evs)
    print("GPU span %.2f ms, busy %.2f ms, idle %.2f ms in %d gaps" % (span / 1e3, busy / 1e3, tot / 1e3, len(gaps)))
    for g, a, b in sorted(gaps, reverse=True)[:30]:
        print("  gap %7.1f us  after %-50s before %s" % (g, a, b))
