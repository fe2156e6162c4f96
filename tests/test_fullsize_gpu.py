"""GPU, BASELINE-size inputs (batch of 8 ScanNet-shaped scenes x 100 000 points): size-independent properties of the
hot path where the CPU oracle would take too long — round trips, symmetry, idempotence, linearity, cardinalities."""
import numpy as np
import pytest
import torch

from fcaf3d_b200 import me as ME
from fcaf3d_b200 import ops
from fcaf3d_b200.scenes import make_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def batch(dev):
    pts = [torch.from_numpy(make_scene("S0", 100 * i)[0]).to(dev) for i in range(8)]
    coords, feats = ops.quantize_points(pts, 0.01)
    return pts, coords, feats


def test_hash_round_trip_and_idempotence(dev, batch):
    pts, coords, feats = batch
    x = ME.SparseTensor(coordinates=coords, features=feats)
    n = len(x)
    assert 8 * 95000 < n < 8 * 100000
    m = x._map()
    # every input point finds its voxel, and the voxel is the one the inverse map names
    rows = ops.map_lookup(m.table, m.cap, coords)
    assert torch.equal(rows.long(), x.inverse_mapping)
    assert torch.equal(x.C[x.inverse_mapping], coords)
    # representatives are the first occurrences: unique_index is increasing and maps back to itself
    assert bool((x.unique_index[1:] > x.unique_index[:-1]).all())
    assert torch.equal(x.inverse_mapping[x.unique_index], torch.arange(n, device=dev))
    # dedup of a deduplicated set is the identity
    y = ME.SparseTensor(coordinates=x.C, features=x.F)
    assert torch.equal(y.C, x.C) and torch.equal(y.unique_index, torch.arange(n, device=dev))
    # batch offsets partition the rows by scene
    off = m.off_host
    assert off[0] == 0 and off[-1] == n and all(bool((x.C[off[b]:off[b + 1], 0] == b).all()) for b in range(8))


def test_stride_maps_and_kernel_map_symmetry(dev, batch):
    _, coords, feats = batch
    x = ME.SparseTensor(coordinates=coords, features=feats)
    cm = x.coordinate_manager
    key = x.coordinate_map_key
    sizes = [len(x)]
    for _ in range(3):
        key = cm.stride(key, 2)
        sizes.append(cm.size(key))
        c = cm.get_coordinates(key)
        ts = cm.get_map(key).ts
        assert bool((c[:, 1:] % ts == 0).all())
    assert sizes == sorted(sizes, reverse=True)
    # stride-1 3^3 kernel map on the stride-8 map: (i -> o at offset k) <=> (o -> i at offset 26 - k)
    km = cm.kernel_map(key, key, 3)
    n = cm.size(key)
    nbr = km.nbr[:, :n]
    assert torch.equal(nbr[13], torch.arange(n, device=dev, dtype=torch.int32))
    for k in (0, 5, 12, 20):
        o = torch.nonzero(nbr[k] >= 0).squeeze(1)
        i = nbr[k][o].long()
        assert torch.equal(nbr[26 - k][i].long(), o)
    # pair count is symmetric too
    assert int((nbr[3] >= 0).sum()) == int((nbr[23] >= 0).sum())
    # children map: 8 rows per parent, parents recovered by flooring
    g = cm.generate_children(key)
    gc = cm.get_coordinates(g)
    assert gc.shape[0] == 8 * n
    ts = cm.get_map(key).ts
    par = gc.clone()
    par[:, 1:] = torch.div(par[:, 1:], ts, rounding_mode="floor") * ts
    assert torch.equal(par.view(n, 8, 4)[:, 0], cm.get_coordinates(key))
    assert bool((par.view(n, 8, 4) == par.view(n, 8, 4)[:, :1]).all())


def test_convolution_is_linear_at_full_size(dev, batch):
    _, coords, feats = batch
    x = ME.SparseTensor(coordinates=coords, features=feats)
    cm = x.coordinate_manager
    key = cm.stride(cm.stride(cm.stride(x.coordinate_map_key, 2), 2), 2)
    n = cm.size(key)
    km = cm.kernel_map(key, key, 3)
    torch.manual_seed(0)
    a = torch.randn(n, 64, device=dev).to(torch.bfloat16)
    b = torch.randn(n, 64, device=dev).to(torch.bfloat16)
    w = torch.randn(27, 64, 64, device=dev) * 0.05
    wt = w.transpose(1, 2).contiguous().to(torch.bfloat16)
    conv = lambda t: ops.sparse_conv(t, w, km, n, wt_bf16=wt, out_dtype=torch.float32)
    ya, yb = conv(a), conv(b)
    s = (a.float() * 0.5 + b.float() * 0.25).to(torch.bfloat16)   # exactly representable combination weights
    ys = conv(s)
    ref = 0.5 * ya + 0.25 * yb
    # the only error is the bf16 rounding of the combined input
    assert ((ys - ref).abs().max() / ref.abs().max()).item() < 1e-2
    # zero input -> exactly zero output; tensor-core and CUDA-core kernels agree on the same operands
    assert float(conv(torch.zeros_like(a)).abs().max()) == 0.0
    ops.set_tensor_cores(False)
    try:
        y_cc = ops.sparse_conv(a, wt.transpose(1, 2).float().contiguous(), km, n, out_dtype=torch.float32)
    finally:
        ops.set_tensor_cores(True)
    assert ((ya - y_cc).abs().max() / y_cc.abs().max()).item() < 5e-4


def test_prune_and_topk_cardinalities(dev, batch):
    _, coords, feats = batch
    x = ME.SparseTensor(coordinates=coords, features=feats)
    m = x._map()
    torch.manual_seed(1)
    scores = torch.randn(len(x), device=dev)
    for k in (1, 1000, 50000, 10 ** 6):
        mask = ops.topk_mask(scores, m.off_dev, 8, k)
        for b in range(8):
            seg = slice(m.off_host[b], m.off_host[b + 1])
            nb = m.off_host[b + 1] - m.off_host[b]
            assert int(mask[seg].sum()) == min(k, nb)
            if 0 < k < nb:  # everything kept is >= everything dropped
                kept, dropped = scores[seg][mask[seg].bool()], scores[seg][~mask[seg].bool()]
                assert float(kept.min()) >= float(dropped.max())
    key, m8, pos, n_keep = x.coordinate_manager.prune(x.coordinate_map_key, ops.topk_mask(scores, m.off_dev, 8, 50000))
    assert n_keep == 8 * 50000 and x.coordinate_manager.get_map(key).off_host == [50000 * b for b in range(9)]
