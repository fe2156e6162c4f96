"""GPU parity (bit-exact, integer work): coordinate hash, dedup, stride / children / union / prune maps and kernel
maps through the C ABI vs the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest
import torch

import oracle.me_cpu as OME
from fcaf3d_b200 import me as ME
from fcaf3d_b200 import ops
from util import random_coords, sort_rows, surface_coords

pytestmark = pytest.mark.gpu


def _pair(coords_np, dev, ts=1):
    ct = torch.from_numpy(coords_np)
    f = torch.arange(len(ct), dtype=torch.float32)[:, None].repeat(1, 2)
    return (ME.SparseTensor(f.to(dev), coordinates=ct.to(dev), tensor_stride=ts),
            OME.SparseTensor(f, coordinates=ct, tensor_stride=ts))


@pytest.mark.parametrize("n,batch,seed", [(1, 1, 0), (50, 1, 1), (5000, 3, 2), (40000, 8, 3)])
def test_insert_and_map_bit_exact(dev, n, batch, seed):
    c = random_coords(n, batch, extent=12 if n > 100 else 3, seed=seed, dup_frac=0.3)
    x, o = _pair(c, dev)
    assert torch.equal(x.C.cpu(), o.C)                      # same rows, same (first-occurrence) order
    assert torch.equal(x.unique_index.cpu(), o.unique_index)
    assert torch.equal(x.inverse_mapping.cpu(), o.inverse_mapping)
    assert torch.equal(x.F.cpu(), o.F)
    # the hash table finds every row and rejects absent coordinates
    m = x._map()
    rows = ops.map_lookup(m.table, m.cap, x.C)
    assert torch.equal(rows.cpu(), torch.arange(len(x), dtype=torch.int32))
    far = x.C.clone()
    far[:, 1] += 100000
    assert bool((ops.map_lookup(m.table, m.cap, far) == -1).all())


def test_ungrouped_batches_and_range_errors(dev):
    c = random_coords(300, 3, extent=5, seed=4)
    perm = np.random.default_rng(0).permutation(len(c))
    x, o = _pair(c[perm], dev)
    xs, = sort_rows(x.C)
    os_, = sort_rows(o.C)
    assert torch.equal(xs, os_)
    assert torch.equal(x.C.cpu()[x.inverse_mapping.cpu()], torch.from_numpy(c[perm]))
    bad = torch.tensor([[0, 0, 0, 0], [0, 1 << 20, 0, 0]], dtype=torch.int32, device=dev)
    with pytest.raises(Exception, match="packed key range"):
        ME.SparseTensor(torch.zeros(2, 1, device=dev), coordinates=bad)


def test_empty_and_single(dev):
    x = ME.SparseTensor(torch.zeros(1, 3, device=dev), coordinates=torch.tensor([[0, -7, 3, 2]], dtype=torch.int32,
                                                                                 device=dev))
    assert len(x) == 1 and x.C.tolist() == [[0, -7, 3, 2]]
    pos, total = ops.scan_mask(torch.zeros(0, dtype=torch.uint8, device=dev))
    assert total.item() == 0


@pytest.mark.parametrize("n", [1, 4095, 4096, 4097, 300001])
def test_scan_matches_cumsum(dev, n):
    g = torch.Generator().manual_seed(n)
    mask = (torch.rand(n, generator=g) < 0.37).to(torch.uint8).to(dev)
    pos, total = ops.scan_mask(mask)
    ref = torch.cumsum(mask.long(), 0) - mask.long()
    assert torch.equal(pos.long(), ref) and total.item() == int(mask.sum())


def test_stride_children_union_prune_bit_exact(dev):
    c = surface_coords(3000, batch=3, seed=1)
    x, o = _pair(c, dev)
    cm, ocm = x.coordinate_manager, o.coordinate_manager
    k2, ok2 = cm.stride(x.coordinate_map_key, 2), ocm.stride(o.coordinate_map_key, 2)
    k4, ok4 = cm.stride(k2, 2), ocm.stride(ok2, 2)
    k8, ok8 = cm.stride(k4, 2), ocm.stride(ok4, 2)
    for k, ok in ((k2, ok2), (k4, ok4), (k8, ok8)):
        assert torch.equal(cm.get_coordinates(k).cpu(), torch.from_numpy(ocm.get_map(ok).coords))
        assert cm.get_map(k).off_host == ocm.get_map(ok).batch_offsets(3).tolist()
    assert cm.stride(x.coordinate_map_key, 2) is k2                      # cached (ME A.3)
    g, og = cm.generate_children(k8), ocm.generate_children(ok8)
    assert torch.equal(cm.get_coordinates(g).cpu(), torch.from_numpy(ocm.get_map(og).coords))
    assert cm.get_map(g).ts == 4
    # union of the backbone map (stride 4) with the generated one
    u, dst_a, dst_b = cm.union(k4, g)
    ou, odst_a, odst_b = ocm.union(ok4, og)
    assert torch.equal(cm.get_coordinates(u).cpu(), torch.from_numpy(ocm.get_map(ou).coords))
    assert torch.equal(dst_a.cpu().long(), odst_a) and torch.equal(dst_b.cpu().long(), odst_b)
    # prune with a random mask keeps order
    mask = torch.rand(cm.size(u), generator=torch.Generator().manual_seed(3)) < 0.4
    pk, _, _, n_keep = cm.prune(u, mask.to(dev))
    opk = ocm.prune(ou, mask)
    assert n_keep == int(mask.sum())
    assert torch.equal(cm.get_coordinates(pk).cpu(), torch.from_numpy(ocm.get_map(opk).coords))


@pytest.mark.parametrize("ksize,stride", [(3, 1), (3, 2), (2, 2), (1, 2)])
def test_kernel_map_bit_exact(dev, ksize, stride):
    c = surface_coords(3000, batch=2, seed=5)
    x, o = _pair(c, dev)
    cm, ocm = x.coordinate_manager, o.coordinate_manager
    ik, oik = x.coordinate_map_key, o.coordinate_map_key
    ok_, ook = (ik, oik) if stride == 1 else (cm.stride(ik, stride), ocm.stride(oik, stride))
    km = cm.kernel_map(ik, ok_, ksize)
    ref = ocm.neighbour_table(oik, ook, ksize)
    n_out = cm.size(ok_)
    assert km.nbr.shape == (ksize ** 3, (n_out + 127) // 128 * 128)
    assert torch.equal(km.nbr[:, :n_out].cpu().long(), torch.from_numpy(ref))
    assert bool((km.nbr[:, n_out:] == -1).all())
    tiles = (n_out + 127) // 128
    pad = np.full((ksize ** 3, tiles * 128), -1)
    pad[:, :n_out] = ref
    want = ((pad.reshape(ksize ** 3, tiles, 128) >= 0).any(2) * (1 << np.arange(ksize ** 3))[:, None]).sum(0)
    assert km.tile_mask.cpu().numpy().astype(np.int64).tolist() == want.tolist()
    if ksize == 3 and stride == 1:
        assert torch.equal(km.nbr[13, :n_out].cpu(), torch.arange(n_out, dtype=torch.int32))  # centre = identity


@pytest.mark.parametrize("ksize", [2, 3])
def test_strided_kernel_map_from_the_input_side_equals_hashed(dev, ksize, monkeypatch):
    """Stride-2 tables built from the input rows (probes of the OUT table) against the per-output-row hash probes:
    same table, same tile masks, at tensor strides 1 -> 2 and 2 -> 4, negative coordinates included."""
    monkeypatch.setattr(ME, "STRIDED_FROM_INPUT", 2)                  # every stride-2 map, whatever its in/out ratio
    c = np.concatenate([surface_coords(3000, batch=2, seed=7), random_coords(2000, batch=3, extent=30, seed=9)[4000:]])
    x, _ = _pair(c, dev)
    cm = x.coordinate_manager
    k1 = x.coordinate_map_key
    k2 = cm.stride(k1, 2)
    k4 = cm.stride(k2, 2)
    for ik, ok_ in ((k1, k2), (k2, k4)):
        src, dst = cm.get_map(ik), cm.get_map(ok_)
        if ksize == 3:
            km = cm.kernel_map(ik, ok_, ksize)                      # input-driven path
        else:                                                       # kernel 2 (pooling) is not routed there: call it directly
            t2, c2 = dst.ensure_table()
            km = ops.kernel_map_strided(t2, c2, src.coords, dst.n, 2, src.ts)
        table, cap = src.ensure_table()
        hashed = ops.kernel_map(table, cap, dst.coords, ksize, src.ts)
        assert torch.equal(km.nbr, hashed.nbr) and torch.equal(km.tile_mask, hashed.tile_mask)
        assert int((km.nbr >= 0).sum()) >= src.n                    # every input row feeds at least its own parent


def test_children_kernel_map_derived_equals_hashed(dev):
    """3^3 table of a generated (all-children) map: the table derived from the parent's table must equal the
    hash-probed one and the oracle's; the derived tile mask is conservative (all 27 offsets)."""
    c = surface_coords(4000, batch=3, seed=11)
    x, o = _pair(c, dev)
    cm, ocm = x.coordinate_manager, o.coordinate_manager
    k4, ok4 = cm.stride(cm.stride(x.coordinate_map_key, 2), 2), ocm.stride(ocm.stride(o.coordinate_map_key, 2), 2)
    g, og = cm.generate_children(k4), ocm.generate_children(ok4)
    km = cm.kernel_map(g, g, 3)                                    # derived path (children_of is set)
    n = cm.size(g)
    assert n == 8 * cm.size(k4) and km.n_out == n
    gm = cm.get_map(g)
    table, cap = gm.ensure_table()
    hashed = ops.kernel_map(table, cap, gm.coords, 3, gm.ts)
    assert torch.equal(km.nbr[:, :n], hashed.nbr[:, :n])
    assert bool((km.nbr[:, n:] == -1).all())
    assert torch.equal(km.nbr[:, :n].cpu().long(), torch.from_numpy(ocm.neighbour_table(og, og, 3)))
    assert bool(((km.tile_mask & hashed.tile_mask) == hashed.tile_mask).all())
    # dilation 2 or another kernel size on the same map still takes the hash path
    km2 = cm.kernel_map(g, g, 3, 2)
    assert torch.equal(km2.nbr[:, :n].cpu().long(), torch.from_numpy(ocm.neighbour_table(og, og, 3, 2)))


def test_prefetched_maps_equal_the_lazy_ones(dev):
    """stride_prefetch / union_prefetch / the raw-row build of the first strided map (asynchronous row counts) give the
    same maps, bit for bit, as the lazy calls; prune with host-known keep counts equals prune with a device read."""
    c = surface_coords(6000, batch=3, seed=21)
    c = torch.from_numpy(np.concatenate([c, c[::7]], 0))           # duplicates: the input map really deduplicates
    c = c[torch.sort(c[:, 0], stable=True).indices].contiguous().to(dev)
    lazy = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c.clone())
    hinted = c.clone()
    hinted._b2_num_batches, hinted._b2_prefetch_stride = 3, 2
    pre = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=hinted)
    cl, cp = lazy.coordinate_manager, pre.coordinate_manager
    assert torch.equal(lazy.C, pre.C) and len(cp._pending) == 1
    kl, kp = lazy.coordinate_map_key, pre.coordinate_map_key
    for _ in range(3):
        nl = cl.stride(kl, 2)
        cp.stride_prefetch(kp, 2)
        np_ = cp.stride(kp, 2)
        assert torch.equal(cl.get_coordinates(nl), cp.get_coordinates(np_))
        assert cl.get_map(nl).off_host == cp.get_map(np_).off_host
        kl, kp = nl, np_
    assert not cp._pending
    # union of the stride-4 map with the children of the stride-8 map, prefetched vs lazy
    l4, p4 = cl.stride(cl.stride(lazy.coordinate_map_key, 2), 2), cp.stride(cp.stride(pre.coordinate_map_key, 2), 2)
    gl, gp = cl.generate_children(kl), cp.generate_children(kp)
    cp.union_prefetch(p4, gp)
    ul, al, bl = cl.union(l4, gl)
    up, ap, bp = cp.union(p4, gp)
    assert torch.equal(cl.get_coordinates(ul), cp.get_coordinates(up)) and torch.equal(al, ap) and torch.equal(bl, bp)
    # prune: counts known on the host
    um = cl.get_map(ul)
    mask = ops.topk_mask(torch.rand(um.n, device=dev), um.off_dev, 3, 500)
    keep = [min(um.off_host[b + 1] - um.off_host[b], 500) for b in range(3)]
    k1, _, _, n1 = cl.prune(ul, mask)
    k2, _, _, n2 = cp.prune(up, mask, keep_counts=keep)
    assert n1 == n2 and torch.equal(cl.get_coordinates(k1), cp.get_coordinates(k2))
    assert cl.get_map(k1).off_host == cp.get_map(k2).off_host
