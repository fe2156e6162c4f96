"""Micro-benchmark of the sparse-conv kernels on real FCAF3D layer shapes (batch of 8 S0 scenes).
    python tools/bench_conv.py [level ...]      level in {ts8, ts16, ts32, ts64, neck8}
Prints per (shape) the CUDA-event time, algorithmic TFLOP/s (2*P*Cin*Cout) and effective gather GB/s."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fcaf3d_b200 import me as ME  # noqa: E402
from fcaf3d_b200 import ops  # noqa: E402
from fcaf3d_b200.scenes import make_scene  # noqa: E402

dev = torch.device("cuda", 0)
pts = [torch.from_numpy(make_scene("S0", 100 * i)[0]).to(dev) for i in range(8)]
coords, feats = ops.quantize_points(pts, 0.01)
x = ME.SparseTensor(coordinates=coords, features=feats)
cm = x.coordinate_manager
keys = {1: x.coordinate_map_key}
for ts in (2, 4, 8, 16, 32, 64):
    keys[ts] = cm.stride(keys[ts // 2], 2)
levels = sys.argv[1:] or ["ts8", "ts16", "ts32", "ts64", "neck8"]
shapes = {"ts8": (8, 64, 64), "ts16": (16, 128, 128), "ts32": (32, 256, 256), "ts64": (64, 512, 512)}
iters = int(os.environ.get("ITERS", "20"))
for name in levels:
    if name == "neck8":
        key = cm.generate_children(keys[16])
        cin = cout = 64
    else:
        ts, cin, cout = shapes[name]
        key = keys[ts]
    n = cm.size(key)
    km = cm.kernel_map(key, key, 3)
    pairs = ops.kmap_pairs(km, n)
    torch.manual_seed(0)
    f = torch.randn(n, cin, device=dev).to(torch.bfloat16)
    w = torch.randn(27, cin, cout, device=dev) * 0.05
    wt = w.transpose(1, 2).contiguous().to(torch.bfloat16)
    if ops.SORT_MIN_ROWS and n >= ops.SORT_MIN_ROWS:
        ops.kmap_sorted(km)          # time the steady state: the table is sorted from its second use on
    for _ in range(3):
        y = ops.sparse_conv(f, w, km, n, wt_bf16=wt)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        y = ops.sparse_conv(f, w, km, n, wt_bf16=wt)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / iters
    flops = 2.0 * pairs * cin * cout
    gather = pairs * cin * 2 + (n + 127) // 128 * 27 * cin * cout * 2 * (1 if cout <= 128 else cout // 128)
    print("%-6s rows=%7d pairs=%9d (%.1f/row) C=%d->%d : %8.1f us  %7.1f TFLOP/s  L2->SM %6.0f GB/s" %
          (name, n, pairs, pairs / n, cin, cout, ms * 1000, flops / ms / 1e9, gather / ms / 1e6))
    if os.environ.get("TC_DBG"):
        # timing experiments of a -DB2_TC_DBG build (B2_NVCC_EXTRA=-DB2_TC_DBG python -m fcaf3d_b200.build): see conv_tc.cu
        from fcaf3d_b200._lib import lib
        import ctypes
        raw = ctypes.CDLL(lib()._name) if hasattr(lib(), "_name") else lib()
        for v in [int(t) for t in os.environ["TC_DBG"].split(",")]:
            raw.b2_tc_debug(ctypes.c_int(v))
            if hasattr(raw, "b2_tc_debug_stats"):
                raw.b2_tc_debug_stats((ctypes.c_ulonglong * 16)(), 1)
                raw.b2_tcp_debug_stats((ctypes.c_ulonglong * 16)(), 1)
            for _ in range(2):
                ops.sparse_conv(f, w, km, n, wt_bf16=wt)
            torch.cuda.synchronize()
            s.record()
            for _ in range(iters):
                ops.sparse_conv(f, w, km, n, wt_bf16=wt)
            e.record()
            torch.cuda.synchronize()
            print("%-6s   dbg=%2d : %8.1f us" % (name, v, s.elapsed_time(e) / iters * 1000))
            if hasattr(raw, "b2_tc_debug_stats"):
                st = (ctypes.c_ulonglong * 16)()
                raw.b2_tc_debug_stats(st, 1)
                c = max(1, st[0])
                names = ["ctas", "stages", "prologue", "mma_wait_full", "mma_issue", "prod_wait_empty", "prod_issue",
                         "wt_wait_empty", "accum_wait", "epilogue", "lifetime"]
                if st[0]:
                    print("         per CTA (cycles): " + "  ".join("%s=%.0f" % (names[i], st[i] / c) for i in range(1, 11)))
                raw.b2_tcp_debug_stats(st, 1)
                if st[0]:
                    c = st[0]
                    names = ["ctas", "stages", "tiles", "mma_wait_full", "mma_issue", "mma_wait_acc", "prod_stage_nbr",
                             "prod_wait_empty", "prod_issue", "epi_wait", "epi", "lifetime", "prologue", "lifetime_ns"]
                    print("         persistent, per CTA: " + "  ".join("%s=%.0f" % (names[i], st[i] / c) for i in range(1, 14)) +
                          "  longest_ns=%d longest_cycles=%d" % (st[14], st[15]))
        raw.b2_tc_debug(ctypes.c_int(0))
        continue
    if ops.SORT_MIN_ROWS and km.sorted is not None:
        import copy
        ts_ = []
        for _ in range(3):
            km2 = copy.copy(km)
            km2.sorted = None
            torch.cuda.synchronize()
            s.record()
            ops.kmap_sorted(km2)
            e.record()
            torch.cuda.synchronize()
            ts_.append(s.elapsed_time(e) * 1000)
        tm_s = km.sorted[2].long()
        popc = sum(((tm_s >> b) & 1).sum().item() for b in range(27)) / tm_s.numel()
        print("%-6s   kmap_sort %.1f us; offsets per tile after sorting %.2f" % (name, min(ts_), popc))
    # weight gradient of the same layer (bf16 operands, fp32 dW)
    from fcaf3d_b200.autograd import conv_wgrad  # noqa: E402
    g = torch.randn(n, cout, device=dev).to(torch.bfloat16)
    for _ in range(3):
        dw = conv_wgrad(f, g, km, 27, cin, cout)
    torch.cuda.synchronize()
    s.record()
    for _ in range(iters):
        dw = conv_wgrad(f, g, km, 27, cin, cout)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / iters
    print("%-6s   wgrad (incl. dW memset)                      : %8.1f us  %7.1f TFLOP/s  gathered %6.0f GB/s" %
          (name, ms * 1000, flops / ms / 1e9, pairs * cin * 2 / ms / 1e6))
