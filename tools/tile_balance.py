import os, sys, torch
sys.path.insert(0, '/root/repo')
from fcaf3d_b200 import me as ME, ops
from fcaf3d_b200.scenes import make_scene
dev = torch.device("cuda", 0)
pts = [torch.from_numpy(make_scene("S0", 100 * i)[0]).to(dev) for i in range(8)]
coords, feats = ops.quantize_points(pts, 0.01)
x = ME.SparseTensor(coordinates=coords, features=feats)
cm = x.coordinate_manager
keys = {1: x.coordinate_map_key}
for ts in (2, 4, 8, 16):
    keys[ts] = cm.stride(keys[ts // 2], 2)
for name, key in (("ts8", keys[8]), ("neck8", cm.generate_children(keys[16]))):
    n = cm.size(key)
    km = cm.kernel_map(key, key, 3)
    ops.kmap_sorted(km)
    tm = km.sorted[2].long()
    pc = torch.zeros_like(tm)
    for b in range(27):
        pc += (tm >> b) & 1
    pc = pc.cpu()
    nt = pc.numel()
    G = 296
    for mode in ("rr", "serp"):
        sums = [0] * G
        for idx in range(nt):
            i, c = divmod(idx, G)
            if mode == "serp" and (i & 1):
                c = G - 1 - c
            sums[c] += int(pc[nt - 1 - idx])
        print(name, mode, "tiles", nt, "stages mean %.1f max %d min %d" % (sum(sums) / G, max(sums), min(sums)))
    # greedy dynamic (heavy-first list scheduling)
    import heapq
    h = [(0, c) for c in range(G)]
    for idx in range(nt):
        s, c = heapq.heappop(h)
        heapq.heappush(h, (s + int(pc[nt - 1 - idx]), c))
    print(name, "dynamic max %d" % max(s for s, _ in h))
