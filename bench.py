#!/usr/bin/env python
"""bench.py — FCAF3D hot-path throughput (scenes/s) of fcaf3d_b200 on synthetic ScanNet / SUN RGB-D / S3DIS-shaped scenes.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--dtype bf16|fp32]
                    [--workload scannet|sunrgbd|s3dis1m] [--mode infer|train] [--configs all|none]

Headline (BASELINE.json configs[1]): one "step" = one full inference pass (quantise -> hash -> ME-ResNet34 backbone ->
FCAF3D neck/head -> top-k -> decode -> per-class BEV NMS) over one batch of 8 synthetic ScanNet-shaped scenes (S0 family,
100 000 points each).  N > 1 (torchrun): every rank runs its own batch (scene sharding, no collective) -> "weak" scaling.
The same JSON line carries a `configs` object with the other BASELINE configs measured in the same run - SUN RGB-D-shape
inference with rotated NMS (configs[2]), S3DIS-1M-shape inference (configs[3]), the data-parallel training step on
ScanNet shapes (axis-aligned IoU loss) and on SUN RGB-D shapes (rotated IoU loss) with DDP / NCCL (configs[4]), the fp32
companion of the headline, and the standalone NMS kernel against the reference's own pcdet_nms.
Prints ONE JSON line (rank 0).  See DESIGN.md §Measurement for every field.
"""
import argparse
import atexit
import ctypes
import gc
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (dataset, scene preset, scenes per GPU per step, metric, workload description)
    "scannet": ("scannet", "S0", 8, "ScanNet-shape scenes/sec, FCAF3D full inference incl. axis-aligned BEV NMS, batch 8",
                "FCAF3D ScanNet-18class full inference incl. axis-aligned 3D(BEV) NMS, batch 8 synthetic S0 scenes x 100k pts"),
    "sunrgbd": ("sunrgbd", "S3", 8, "SUN RGB-D-shape scenes/sec, FCAF3D inference with rotated BEV NMS, batch 8",
                "FCAF3D SUN RGB-D-10class rotated boxes (Mobius) inference incl. rotated BEV NMS, batch 8 synthetic S3 scenes "
                "x 100k pts"),
    "s3dis1m": ("s3dis", "S2", 1, "S3DIS-1M-shape scenes/sec, FCAF3D inference, 1 scene per GPU per step",
                "FCAF3D S3DIS-5class large room (1M points) inference, scene-sharded, 1 synthetic S2 scene per GPU"),
}
N_CLS = {"scannet": 18, "sunrgbd": 10, "s3dis": 5}


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=100):
        self.rows, self.proc, self.index, self.period_ms = [], None, index, period_ms
        self.first = 0

    def mark(self):
        """Samples taken before this call (process start-up, warm-up) are not reported."""
        self.first = len(self.rows)

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", str(self.period_ms)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
            self.proc = None

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows[self.first:]:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_forward_baseline(n_scenes=1, threads=None):
    """CPU restatement of the ME path (oracle/model_cpu.py): backbone + neck + head forward of `n_scenes` S0 scenes
    processed one by one (configs[0]; the reference has no CPU NMS: pcdet_nms aborts on CPU tensors)."""
    import torch
    from fcaf3d_b200.scenes import make_scene
    from fcaf3d_b200.testing import deterministic_init
    from oracle import model_cpu
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    det = deterministic_init(model_cpu.Detector("scannet"), 0, cls_bias=0.0).eval()
    scenes = [torch.from_numpy(make_scene("S0", 100 * i)[0]) for i in range(n_scenes)]
    t0 = time.perf_counter()
    for p in scenes:
        det([p])
    dt = time.perf_counter() - t0
    return n_scenes / dt, dt, threads


def run_reference(args):
    """--impl reference: the reference's ME CPU backend cannot be installed (no MinkowskiEngine source, no network; probed
    on the GPU box as well, profiles/r02_me_probe.txt), so the CPU port in oracle/ is timed on this box's host cores
    (kind = "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch  # noqa: F401
    _, _, _, metric, workload = WORKLOADS["scannet"]
    per_step = []
    for i in range(args.warmup + args.steps):
        v, dt, threads = cpu_forward_baseline(1)
        if i >= args.warmup:
            per_step.append(dt)
    ms = 1000.0 * sum(per_step) / len(per_step)
    value = 1000.0 / ms
    line = {"impl": "reference", "metric": metric, "value": value, "unit": "scenes/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload, "sample": "1 S0 scene per step, forward only (no CPU NMS in the reference)"},
            "cpu_baseline": {"value": value, "unit": "scenes/s", "cores": threads, "kind": "port",
                             "sample": "1 S0 scene (100k points) backbone+neck+head forward per step"},
            "e2e": {"value": value, "unit": "scenes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def kernel_digest():
    """sha256 of the tcgen05 convolution source the built library came from (ties profiles/conv_tc_traffic.json to it)"""
    import hashlib
    h = hashlib.sha256()
    for f in ("conv_tc.cu", "conv_tcp.cu", "tc_common.cuh"):
        h.update(open(os.path.join(ROOT, "fcaf3d_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def roofline_from_profile(prof, ms, use_traffic_file, main=None, brief=False):
    """Aggregates ops.PROFILE records [name, start, end, flops, bytes] of ONE step into the bench line's roofline object
    (dominant kernel = the tcgen05 convolution; the other profiled kernels ride along in `other_kernels`)."""
    agg = {}
    for name, s, e, fl, by in prof:
        a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
        a[0] += 1
        a[1] += s.elapsed_time(e)
        a[2] += fl
        a[3] += by
    pk, pk_kind = peaks()
    kernels = {}
    for name, (cnt, t_ms, fl, by) in agg.items():
        kernels[name] = {"launches": cnt, "ms": round(t_ms, 4), "tflops": fl / t_ms / 1e9 if fl else None,
                         "gbs": by / t_ms / 1e6, "hbm_frac": by / t_ms / 1e6 / pk["hbm_gbs"]}
    main_k = main or ("conv_tc" if "conv_tc" in agg else "conv_cc")
    cnt, t_ms, fl, by = agg[main_k]
    peak = pk["bf16_tflops_sustained"] if main_k in ("conv_tc", "wgrad_tc") else 75.0
    traffic = None
    try:  # DRAM bytes per launch of the dominant kernel from the committed ncu capture; refused when the kernel changed
        if main_k == "conv_tc" and use_traffic_file:
            rec = json.load(open(os.path.join(ROOT, "profiles", "conv_tc_traffic.json")))
            if rec.get("kernel_digest") == kernel_digest():
                traffic = rec["dram_bytes_per_launch"]
    except Exception:
        pass
    out = {"kernel": main_k, "bound": "tensor", "achieved": fl / t_ms / 1e9, "peak": peak, "unit": "TFLOP/s",
           "frac": fl / t_ms / 1e9 / peak, "traffic": traffic, "algorithmic_bytes_per_launch": by / cnt,
           "peak_source": pk_kind, "launches_per_step": cnt, "avg_launch_ms": t_ms / cnt, "share_of_step": t_ms / ms,
           "algorithmic_flops_per_step": fl}
    if not brief:
        out["other_kernels"] = kernels
    return out


class Ctx:
    """process-wide state of one bench run"""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)

    def barrier(self):
        import torch.distributed as dist
        if self.world > 1:
            dist.barrier()

    def max_over_ranks(self, ms):
        import torch
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()


def make_model(ctx, workload, calibrate=True):
    """FCAF3D config of the workload, deterministic random weights; cls bias 0 so that candidates pass score_thr (SURVEY
    F8); inference: BatchNorm running statistics calibrated by one train-mode pass (what a trained checkpoint carries)."""
    import torch
    from fcaf3d_b200.detector import build_fcaf3d
    from fcaf3d_b200.scenes import make_scene
    from fcaf3d_b200.testing import deterministic_init
    dataset, preset, batch, _, _ = WORKLOADS[workload]
    det = deterministic_init(build_fcaf3d(dataset), 0, cls_bias=0.0).to(ctx.dev)
    scenes = [make_scene(preset, 100 * (ctx.rank * batch + i), n_classes=N_CLS[dataset]) for i in range(batch)]
    host_pts = [torch.from_numpy(sc[0]).pin_memory() for sc in scenes]
    if calibrate:
        with torch.no_grad():
            det.train()
            for m in det.modules():
                if isinstance(m, torch.nn.BatchNorm1d):
                    m.momentum = 1.0
            det.forward_features_fused([p.to(ctx.dev) for p in host_pts])
            det.eval()
    return det, scenes, host_pts


def bench_infer(ctx, args, workload, dtype_name, steps, warmup, clk=None, brief=False, e2e=True):
    """Inference throughput of one workload: K device-resident steps in one bracket, then the end-to-end loop (pinned host
    inputs copied every step, detections read back every step), then one profiled step for the per-kernel roofline."""
    import torch
    from fcaf3d_b200 import me as ME
    from fcaf3d_b200 import ops
    from fcaf3d_b200.data import ScenePrefetcher
    from fcaf3d_b200.detector import DetectionRead, wait_ready
    dtype = torch.bfloat16 if dtype_name == "bf16" else torch.float32
    ME.set_compute_dtype(dtype)
    ME.set_tf32(dtype_name == "tf32")
    dataset, preset, batch, metric, desc = WORKLOADS[workload]
    det, scenes, host_pts = make_model(ctx, workload)
    dev, world = ctx.dev, ctx.world
    dev_pts = [p.to(dev) for p in host_pts]
    lib = ops._lib.lib()

    def step_device():
        # inputs_ready=True: the points are resident and complete (synchronised before the bracket), so the executor's
        # coordinate stream may start on step i + 1 while step i's convolutions still run
        return det.simple_test_fused(dev_pts, to_host=False, inputs_ready=True)

    if os.environ.get("BENCH_PROFILE"):
        # one step between cudaProfilerStart/Stop for `ncu --profile-from-start off` (numbers printed here are not bench values)
        step_device()
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step_device()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return {"profile_step": "done"}

    for _ in range(max(warmup, 2 * int(getattr(det, "lanes", 1)) + 1)):   # every executor lane sees both of its arenas
        step_device()
    # det.submit: one host thread per executor lane (the executor's call blocks on row-count reads of its coordinate stream;
    # from a single thread the host needed as long to enqueue a step as the GPU to run one).  B2_ASYNC=0: synchronous calls.
    use_submit = os.environ.get("B2_ASYNC", "1") != "0" and getattr(det, "lanes", 1) > 1 and not det.training
    if use_submit:
        for h in [det.submit(dev_pts, read=False, inputs_ready=True) for _ in range(2 * det.lanes)]:
            wait_ready(h.dets())
    torch.cuda.synchronize()
    launches0 = lib.b2_launch_count()
    if clk is not None:
        clk.mark()
    # EXACTLY `steps` steps inside ONE bracket (barrier + synchronize on both sides, CUDA events on the launching stream);
    # L2 is flushed before every step INSIDE the bracket (256 MB write > 126 MB L2: ~45 us per step counted against us),
    # so the host may run ahead of the GPU across steps as it does in a real serving loop.
    gc.collect()
    gc.disable()                # a generation-2 collection (several ms of host time) inside a 100 ms bracket is noise, not signal
    ctx.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    h0 = time.perf_counter()
    recent = []
    for _ in range(steps):
        ctx.flush.fill_(1)
        dets = det.submit(dev_pts, read=False, inputs_ready=True) if use_submit else step_device()
        recent = (recent + [dets])[-4:]
    if use_submit:
        recent = [h.dets() for h in recent]
        dets = recent[-1]
    host_ms = 1000.0 * (time.perf_counter() - h0) / steps      # time the host needed to ENQUEUE a step (incl. its reads)
    for d in recent:            # the last steps' work runs on the executor's own streams: the bracket closes behind them
        wait_ready(d)
    e.record()
    torch.cuda.synchronize()
    gc.enable()
    ctx.barrier()
    ms = ctx.max_over_ranks(s.elapsed_time(e)) / steps
    launches = lib.b2_launch_count() - launches0  # kernels of libfcaf3d_b200.so inside the timed region
    out = {"metric": metric, "workload": desc, "value": world * batch * 1000.0 / ms, "unit": "scenes/s", "ms_per_step": ms,
           "dtype": dtype_name, "scenes_per_gpu": batch, "gpu_launches": launches, "host_ms_per_step": host_ms,
           "detections_per_step": int(dets["count"].item())}
    if e2e:
        # End to end through the public API: every step copies ITS inputs from pinned host memory and reads ITS detections
        # back, all inside one timed bracket; the copy of step i+1 is double-buffered on a side stream
        # (fcaf3d_b200.data.ScenePrefetcher) and overlaps the compute of step i; L2 is flushed inside the loop.
        k_e2e = max(5, steps)
        depth = max(1, int(getattr(det, "lanes", 1)))

        def e2e_loop(k):
            """k steps through the public API, as many reads in flight as the detector keeps steps in flight"""
            pending = []
            for pts in ScenePrefetcher((host_pts for _ in range(k)), dev):
                ctx.flush.fill_(1)
                if use_submit:
                    pending.append(det.submit(pts, read=True))       # lane thread: executor call + async packed read-back
                else:
                    pending.append(DetectionRead(det.simple_test_fused(pts, to_host=False), batch))   # async packed copy
                if len(pending) > depth:
                    r_ = pending.pop(0)             # an earlier step's detections, read while the later ones run
                    r_.result() if use_submit else r_.finish()
            for r_ in pending:
                r_.result() if use_submit else r_.finish()

        e2e_loop(2 * depth + 3)     # warm-up through the SAME path (the pinned staging pool reaches its steady size here)
        ctx.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        e2e_loop(k_e2e)
        e.record()
        torch.cuda.synchronize()
        ms_e2e = ctx.max_over_ranks(s.elapsed_time(e)) / k_e2e
        # bytes actually copied per step by DetectionRead: the count + the first CAP packed rows (box, score, label, scene)
        d2h = 8 + min(DetectionRead.CAP, int(dets["boxes"].shape[0])) * (int(dets["box_dim"]) + 3) * 4
        out["e2e"] = {"value": world * batch * 1000.0 / ms_e2e, "unit": "scenes/s",
                      "h2d_bytes_per_step": sum(p.numel() * 4 for p in host_pts), "d2h_bytes_per_step": d2h,
                      "ms_per_step": ms_e2e}
    if clk is not None:
        out["clocks"] = clk.summary()
    # per-kernel roofline pass (one extra step with CUDA events around the profiled launches)
    ops.PROFILE = []
    step_device()
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    out["roofline"] = roofline_from_profile(prof, ms, workload == "scannet" and dtype_name == "bf16", brief=brief)
    ME.set_compute_dtype(torch.float32)
    ME.set_tf32(False)
    return out


def bench_train(ctx, args, workload, dtype_name, steps, warmup, brief=False):
    """configs[4]: data-parallel training step (8 scenes per GPU): forward, backward, NCCL gradient all-reduce through
    DDP, gradient clipping, AdamW.  At N > 1 also: `ddp_check` (the reduced gradient equals the mean of the ranks' local
    gradients), the all-reduce volume, a standalone NCCL all-reduce of that volume (bus bandwidth) and the all-reduce
    time that stays exposed per step (DDP step time minus the same step without gradient synchronisation)."""
    import contextlib
    import torch
    import torch.distributed as dist
    from fcaf3d_b200 import me as ME
    from fcaf3d_b200 import ops
    from fcaf3d_b200.train import TrainStepModule, make_optimizer, train_step, wrap_ddp
    dtype = torch.bfloat16 if dtype_name == "bf16" else torch.float32
    ME.set_compute_dtype(dtype)
    dataset, preset, batch, metric, desc = WORKLOADS[workload]
    det, scenes, _ = make_model(ctx, workload, calibrate=False)
    dev, world = ctx.dev, ctx.world
    det.train()
    module = wrap_ddp(TrainStepModule(det), dev)
    opt = make_optimizer(module)
    host = [(torch.from_numpy(p).pin_memory(), torch.from_numpy(g), torch.from_numpy(l)) for p, g, l in scenes]
    gtb = [g.to(dev) for _, g, _ in host]
    gtl = [l.to(dev) for _, _, l in host]
    dev_pts = [p.to(dev) for p, _, _ in host]

    def step(_e2e=False):
        total, _ = train_step(module, opt, dev_pts, gtb, gtl)
        return total

    def fetch(e2e):
        return [p.to(dev, non_blocking=True) for p, _, _ in host] if e2e else dev_pts

    loss_ring = [torch.empty((1,), dtype=torch.float32).pin_memory() for _ in range(2)]   # pinned loss read-back

    def run_steps(e2e, k):
        """K training steps, software-pipelined: the coordinate / kernel maps of batch i+1 (and, end to end, its
        host->device copy) are issued between the forward and the backward of batch i (train_step(next_points=...)), so
        the next forward is enqueued while the GPU still runs this backward.  Every step builds its maps exactly once,
        copies its own inputs and (end to end) has its loss read back - one step late, through pinned memory."""
        prepared = det.prepare(fetch(e2e))
        pending, last = None, None
        for i in range(k):
            nxt = fetch(e2e) if i + 1 < k else None
            out = train_step(module, opt, prepared, gtb, gtl, next_points=nxt)
            prepared = out[2] if nxt is not None else None
            if e2e:
                if pending is not None:
                    pending[1].synchronize()
                    last = float(pending[0][0])
                hostbuf = loss_ring[i % 2]
                hostbuf.copy_(out[0].reshape(1), non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                pending = (hostbuf, ev)
            else:
                last = out[0]
        if pending is not None:
            pending[1].synchronize()
            last = float(pending[0][0])
        return last

    host_ms = [None]

    def timed(e2e, k, warm, nosync=False):
        guard = module.no_sync() if (nosync and world > 1) else contextlib.nullcontext()
        with guard:
            if warm:
                run_steps(e2e, warm)
            ctx.barrier()
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            t0 = time.perf_counter()
            out = run_steps(e2e, k)
            host_ms[0] = (time.perf_counter() - t0) * 1e3 / k   # host time to enqueue a step (incl. its blocking reads)
            e.record()
            torch.cuda.synchronize()
        return ctx.max_over_ranks(s.elapsed_time(e)) / k, out

    if os.environ.get("BENCH_PROFILE"):
        step(), step()
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return {"profile_step": "done"}
    lib = ops._lib.lib()
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    launches0 = lib.b2_launch_count()
    ms, _ = timed(False, steps, 0)
    host_enqueue_ms = host_ms[0]
    launches = lib.b2_launch_count() - launches0   # kernels of libfcaf3d_b200.so only (torch's own are not counted)
    ms_e2e, loss = timed(True, max(2, steps // 2), 1)
    n = len(scenes)
    out = {"metric": metric.replace("full inference incl. axis-aligned BEV NMS", "training step").replace(
               "inference with rotated BEV NMS", "training step"),
           "workload": desc.replace(" incl. axis-aligned 3D(BEV) NMS", "").replace(" incl. rotated BEV NMS", "").replace(
               "inference", "training step (fwd+bwd+allreduce+clip+AdamW)"),
           "value": world * n * 1000.0 / ms, "unit": "scenes/s", "ms_per_step": ms, "dtype": dtype_name,
           "scenes_per_gpu": n, "gpu_launches": launches, "host_ms_per_step": host_enqueue_ms,
           "loss": "focal + centerness BCE + %s IoU" % ("rotated" if det.neck_with_head.loss_bbox.with_yaw else "axis-aligned"),
           "optimizer": "AdamW lr 1e-3 wd 1e-4, clip 10", "last_loss": loss,
           "e2e": {"value": world * n * 1000.0 / ms_e2e, "unit": "scenes/s",
                   "h2d_bytes_per_step": sum(p.numel() * 4 for p, _, _ in host), "d2h_bytes_per_step": 4,
                   "ms_per_step": ms_e2e}}
    grad_bytes = sum(p.numel() * 4 for p in module.parameters() if p.requires_grad)
    out["allreduce_bytes_per_step"] = grad_bytes if world > 1 else 0
    if world > 1:
        # (1) gradient equality: local gradients (no_sync) gathered from every rank, their mean against what DDP left in .grad
        params = [p for _, p in sorted(module.named_parameters())]
        picks = [params[i] for i in sorted({0, 1, len(params) // 3, len(params) // 2, 2 * len(params) // 3, len(params) - 2,
                                            len(params) - 1})]

        def sample():
            return torch.cat([p.grad.detach().reshape(-1)[:4096].float() for p in picks])

        opt.zero_grad(set_to_none=True)
        with module.no_sync():
            total, _ = module(dev_pts, gtb, gtl)
            total.backward()
        local = sample()
        gathered = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(gathered, local)
        mean = torch.stack(gathered).mean(0)
        opt.zero_grad(set_to_none=True)
        total, _ = module(dev_pts, gtb, gtl)
        total.backward()
        reduced = sample()
        err = ((reduced - mean).abs().max() / mean.abs().max().clamp(min=1e-30)).item()
        spread = ((gathered[0] - gathered[-1]).abs().max() / mean.abs().max().clamp(min=1e-30)).item()
        opt.zero_grad(set_to_none=True)
        out["ddp_check"] = {"max_rel_err_reduced_vs_mean_of_local": err, "ok": bool(err < 1e-4), "sampled_values": int(local.numel()),
                            "rank_spread": spread, "tolerance": 1e-4}
        # (2) the same volume through one NCCL all-reduce, timed alone
        flat = torch.zeros(grad_bytes // 4, dtype=torch.float32, device=dev)
        for _ in range(2):
            dist.all_reduce(flat)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            dist.all_reduce(flat)
        e.record()
        torch.cuda.synchronize()
        ar_ms = ctx.max_over_ranks(s.elapsed_time(e)) / 5
        del flat
        # (3) what stays exposed: the same steps without gradient synchronisation
        ms_nosync, _ = timed(False, steps, 1, nosync=True)
        out["allreduce"] = {"standalone_ms": ar_ms, "busbw_gbs": 2.0 * (world - 1) / world * grad_bytes / ar_ms / 1e6,
                            "exposed_ms_per_step": ms - ms_nosync, "ms_per_step_without_sync": ms_nosync,
                            "bucket_mb": int(os.environ.get("B2_DDP_BUCKET_MB", "100"))}
    if dtype_name == "bf16":
        ops.PROFILE = []
        step()
        torch.cuda.synchronize()
        prof, ops.PROFILE = ops.PROFILE, None
        out["roofline"] = roofline_from_profile(prof, ms, False, brief=brief)   # conv_tc = forward + input-gradient convolutions
    ME.set_compute_dtype(torch.float32)
    del module, opt, det
    torch.cuda.empty_cache()
    return out


def nms_vs_reference(ctx, n_per_class=(1000, 4000), n_cls=18, n_scenes=8, thr=0.5, reps=5):
    """Standalone NMS: N boxes per class x 18 classes x 8 scenes, random boxes.  The product runs ALL 144
    segments in one batched call (b2_nms_greedy_batched: segment sort + greedy kernel, no host round trip); the baseline
    is the reference's OWN pcdet_nms - its CUDA kernels and its host code iou3d_nms.cpp:90-186 compiled unmodified into
    oracle/_ref/libpcdet_ref_cuda.so - driven the way pcdet_nms_utils.py:81-112 and _nms (fcaf3d_neck_with_head.py:332-374)
    drive it: per (scene, class) a torch sort, a gather, cudaMalloc + mask kernel + blocking D2H + host loop + cudaFree.
    Keep-sets of both sides are compared segment by segment.  The reference library is the yardstick here, exactly like the
    CPU baseline; nothing of it is on the product path."""
    import numpy as np
    import torch
    from fcaf3d_b200 import ops
    from fcaf3d_b200.scenes import random_boxes
    path = os.path.join(ROOT, "oracle", "_ref", "libpcdet_ref_cuda.so")
    if not os.path.exists(path):
        return {"unavailable": "oracle/_ref/libpcdet_ref_cuda.so not built (run __graft_entry__.build() where /root/reference exists)"}
    ref = ctypes.CDLL(path)
    if not hasattr(ref, "ref_nms_gpu_call"):
        return {"unavailable": "oracle/_ref/libpcdet_ref_cuda.so predates ref_nms_gpu_call: rebuild the oracle"}
    ref.ref_nms_gpu_call.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_float, ctypes.c_int]
    ref.ref_nms_gpu_call.restype = ctypes.c_int
    dev = ctx.dev
    out = {}
    for rotated in (False, True):
        for n in n_per_class:
            nseg = n_cls * n_scenes
            segs = [random_boxes(n, 1000 * int(rotated) + s, rotated) for s in range(nseg)]
            boxes = torch.from_numpy(np.concatenate([b for b, _ in segs])).to(dev)
            scores = torch.from_numpy(np.concatenate([s for _, s in segs])).to(dev)
            seg_off = [i * n for i in range(nseg + 1)]
            # ---- product: one batched call
            for _ in range(2):
                keep, order, _, counts = ops.nms_batched(boxes, scores, seg_off, thr, rotated, sort=True)
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(reps):
                keep, order, _, counts = ops.nms_batched(boxes, scores, seg_off, thr, rotated, sort=True)
            e.record()
            torch.cuda.synchronize()
            ours_ms = s.elapsed_time(e) / reps
            # ---- reference: per segment, the reference's own call sequence
            def ref_segment(i):
                b, sc = boxes[i * n:(i + 1) * n], scores[i * n:(i + 1) * n]
                od = sc.sort(0, descending=True)[1]
                bs = b[od].contiguous()
                kp = torch.LongTensor(bs.size(0))
                num = ref.ref_nms_gpu_call(bs.data_ptr(), bs.size(0), kp.data_ptr(), thr, int(rotated))
                return od[kp[:num].cuda()].contiguous()
            ref_segment(0)
            ref_ms = None
            for _ in range(2):       # best of two passes (the first cudaMalloc after a large free can take a second)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                ref_keep = [ref_segment(i) for i in range(nseg)]
                torch.cuda.synchronize()
                dt_ms = 1000.0 * (time.perf_counter() - t0)
                ref_ms = dt_ms if ref_ms is None else min(ref_ms, dt_ms)
            # ---- same keep-sets
            keep_h, order_h = keep.cpu().numpy().astype(bool), order.cpu().numpy()
            same = True
            for i in range(nseg):
                mine = np.sort(order_h[i * n:(i + 1) * n][keep_h[i * n:(i + 1) * n]])
                same = same and np.array_equal(mine, np.sort(ref_keep[i].cpu().numpy()))
            pairs = nseg * n * (n - 1) / 2
            out["%s_%d" % ("rotated" if rotated else "aligned", n)] = {
                "segments": nseg, "boxes_per_segment": n, "ours_ms": ours_ms, "reference_ms": ref_ms,
                "speedup": ref_ms / ours_ms, "keep_sets_equal": bool(same), "kept": int(keep.sum().item()),
                "ours_gpairs_per_s": pairs / ours_ms / 1e6}
    out["reference"] = "pcdet_nms kernels + host code of the reference, unmodified, compiled for sm_100a (oracle/_ref)"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32", "tf32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="scannet", choices=list(WORKLOADS),
                    help="headline of this run: scannet = BASELINE configs[1] (default); sunrgbd = configs[2] rotated boxes + "
                         "rotated NMS; s3dis1m = configs[3] ~1M-point rooms, one scene per step per GPU")
    ap.add_argument("--mode", default="infer", choices=["infer", "train"],
                    help="train = configs[4]: fwd + bwd + NCCL all-reduce + clip + AdamW on 8 scenes per GPU")
    ap.add_argument("--configs", default=None, choices=["all", "none"],
                    help="all = also measure the other BASELINE configs into the line's `configs` object (default for the "
                         "default headline), none = headline only")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    default_headline = args.workload == "scannet" and args.mode == "infer" and args.dtype == "bf16"
    want_configs = (args.configs or ("all" if default_headline else "none")) == "all"

    import torch.distributed as dist
    ctx = Ctx()
    dataset, preset, batch, metric, desc = WORKLOADS[args.workload]
    clk = ClockSampler(ctx.local, period_ms=50)
    if not os.environ.get("BENCH_PROFILE"):
        # started before the warm-up (nvidia-smi needs ~0.1 s to deliver its first line); samples are counted from the first
        # timed bracket of the headline to the end of its end-to-end bracket
        clk.__enter__()
        atexit.register(clk.__exit__)      # the sampler process must not outlive a failed run
    if args.mode == "train":
        head = bench_train(ctx, args, args.workload, args.dtype, args.steps, args.warmup)
        head.setdefault("clocks", clk.summary())
    else:
        head = bench_infer(ctx, args, args.workload, args.dtype, args.steps, args.warmup, clk=clk)
    if "profile_step" in head:
        print(json.dumps(head))
        return
    line = {"metric": head["metric"], "value": head["value"], "unit": "scenes/s", "n_gpus": ctx.world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": head["workload"], "scenes_per_gpu": head["scenes_per_gpu"], "points_per_scene":
                       1000000 if args.workload == "s3dis1m" else 100000, "voxel_size": 0.01,
                       "weights": "deterministic random init, cls bias 0 (all candidates pass score_thr), BN stats "
                                  "calibrated on the batch",
                       "l2": "256 MB flush write before every step, inside the one timed bracket over all steps"
                             if args.mode == "infer" else "activations of one step (GBs) exceed L2",
                       "pipelining": "steps overlap: executor lanes (one host thread each, detector.submit) x (coordinate, "
                                     "feature, tail) streams; python gc is off inside the device-resident bracket only "
                                     "(on in the end-to-end bracket); host_ms_per_step = the submitting thread's time per step, which includes waiting for a free lane slot "
                                     "(back-pressure: two steps per lane outstanding at most)"
                                     if args.mode == "infer" else "maps of batch i+1 built between forward and backward of i",
                       "sharding": "whole scenes per GPU, no collective" if args.mode == "infer" else
                                   "data parallel, 8 scenes per GPU, DDP gradient all-reduce over NCCL"},
            "clocks": head.get("clocks"), "e2e": head["e2e"], "gpu_launches": head["gpu_launches"],
            "roofline": head.get("roofline")}
    for k in ("detections_per_step", "loss", "optimizer", "last_loss"):
        if k in head:
            line["config"][k] = head[k]
    for k in ("host_ms_per_step", "ddp_check", "allreduce", "allreduce_bytes_per_step"):
        if k in head:
            line[k] = head[k]
    if want_configs:
        k_sub = max(5, min(args.steps, 20))
        cfgs = {}
        cfgs["scannet_infer_fp32"] = bench_infer(ctx, args, "scannet", "fp32", max(3, k_sub // 2), 3, brief=True)
        cfgs["scannet_infer_tf32"] = bench_infer(ctx, args, "scannet", "tf32", k_sub, 3, brief=True)
        cfgs["sunrgbd_infer"] = bench_infer(ctx, args, "sunrgbd", "bf16", k_sub, 3, brief=True)
        cfgs["s3dis1m_infer"] = bench_infer(ctx, args, "s3dis1m", "bf16", k_sub, 3, brief=True)
        cfgs["train_scannet"] = bench_train(ctx, args, "scannet", "bf16", k_sub, 3, brief=True)
        cfgs["train_sunrgbd"] = bench_train(ctx, args, "sunrgbd", "bf16", k_sub, 3, brief=True)
        if ctx.rank == 0:
            cfgs["nms_vs_reference"] = nms_vs_reference(ctx)
        ctx.barrier()
        line["configs"] = cfgs
        line["fp32"] = {"value": cfgs["scannet_infer_tf32"]["value"], "unit": "scenes/s",
                        "ms_per_step": cfgs["scannet_infer_tf32"]["ms_per_step"],
                        "note": "same workload with fp32 storage; wide convolutions on the tensor cores as TF32 operands "
                                "(fp32 accumulation, 1e-3 of max per layer); exact fp32 on CUDA cores (the 2e-5 parity "
                                "mode) is configs.scannet_infer_fp32",
                        "exact_fp32_value": cfgs["scannet_infer_fp32"]["value"]}
    clk.__exit__()
    if ctx.rank == 0 and ctx.world == 1 and not args.no_cpu_baseline and args.mode == "infer":
        v, dt, threads = cpu_forward_baseline(1)
        line["cpu_baseline"] = {"value": v, "unit": "scenes/s", "cores": threads, "kind": "port",
                                "sample": "1 S0 scene (100k points), backbone+neck+head forward on the CPU restatement "
                                          "of the ME path (MinkowskiEngine itself unavailable); %.1f s" % dt}
    if ctx.rank == 0:
        print(json.dumps(line))
    if ctx.world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
