"""GPU parity of the head tail: interpolation, top-k masks (incl. ties), head post-processing, box decoding
(vs golden vectors from the reference's own _bbox_pred_to_bbox), candidate assembly."""
import os

import numpy as np
import pytest
import torch

import oracle.me_cpu as OME
from conftest import GOLDEN
from fcaf3d_b200 import me as ME
from fcaf3d_b200 import ops
from util import surface_coords

pytestmark = pytest.mark.gpu


def test_interpolation_matches_oracle(dev):
    c = torch.from_numpy(surface_coords(3000, 2, 7))
    o0 = OME.SparseTensor(torch.zeros(len(c), 1), coordinates=c)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c.to(dev))
    cm, ocm = x0.coordinate_manager, o0.coordinate_manager
    k, ok = cm.stride(cm.stride(x0.coordinate_map_key, 2), 2), ocm.stride(ocm.stride(o0.coordinate_map_key, 2), 2)
    k2, ok2 = cm.stride(k, 2), ocm.stride(ok, 2)
    torch.manual_seed(0)
    f = torch.randn(cm.size(k2), 3)
    src = ME.SparseTensor(f.to(dev), coordinate_map_key=k2, coordinate_manager=cm)
    osrc = OME.SparseTensor(f, coordinate_map_key=ok2, coordinate_manager=ocm)
    q = cm.get_coordinates(k).float()
    got = src.features_at_coordinates(q).cpu()
    want = osrc.features_at_coordinates(q.cpu())
    assert torch.allclose(got, want, atol=1e-6)
    qf = q + torch.rand_like(q) * torch.tensor([0, 3, 3, 3.], device=dev)   # genuinely fractional queries
    assert torch.allclose(src.features_at_coordinates(qf).cpu(), osrc.features_at_coordinates(qf.cpu()), atol=1e-5)


@pytest.mark.parametrize("n,k", [(10, 3), (5000, 1000), (200000, 100000), (7, 7), (7, 100), (300, 0)])
def test_topk_mask_exact_with_ties(dev, n, k):
    g = torch.Generator().manual_seed(n + k)
    segs = [n, max(n // 3, 1), 0, n]
    off = np.concatenate([[0], np.cumsum(segs)]).astype(np.int32)
    vals = torch.randn(int(off[-1]), generator=g)
    vals = torch.round(vals * 8) / 8                     # heavy ties, also at the k-th value
    vals[::97] = float("-inf")
    mask = ops.topk_mask(vals.to(dev), torch.from_numpy(off).to(dev), len(segs), k).cpu().bool()
    for s in range(len(segs)):
        v = vals[off[s]:off[s + 1]]
        kk = min(k, len(v))
        want = torch.zeros(len(v), dtype=torch.bool)
        want[torch.argsort(-v, stable=True)[:kk]] = True  # lowest row first among equals
        assert torch.equal(mask[off[s]:off[s + 1]], want), (s, int(want.sum()), int(mask[off[s]:off[s + 1]].sum()))


def test_decode_boxes_matches_reference_golden(dev):
    g = np.load(os.path.join(GOLDEN, "ops_ref.npz"))
    pts = torch.from_numpy(g["dec_points"]).to(dev)
    p8 = torch.from_numpy(g["dec_pred8"]).to(dev)
    for mode, key, pred in (("fcaf3d", "dec_fcaf3d", p8), ("sin-cos", "dec_sincos", p8),
                            ("naive", "dec_naive", p8[:, :7].contiguous()), ("none", "dec_aligned", p8[:, :6].contiguous())):
        out = ops.decode_boxes(pts, pred, ops.YAW_MODES[mode]).cpu().numpy()
        want = g[key]
        assert np.allclose(out[:, :want.shape[1]], want, rtol=2e-6, atol=2e-6), mode


def test_head_post(dev):
    torch.manual_seed(1)
    n, n_reg, n_cls = 777, 8, 10
    feat = torch.randn(n, 1 + n_reg + n_cls, device=dev)
    scale = torch.tensor([1.3], device=dev)
    ctr, bbox, cls, prune, mx = ops.head_post(feat, n_reg, n_cls, scale, True)
    f = feat.cpu()
    assert torch.equal(ctr.cpu(), f[:, :1]) and torch.equal(cls.cpu(), f[:, 1 + n_reg:])
    assert torch.allclose(bbox.cpu()[:, :6], torch.exp(1.3 * f[:, 1:7]), rtol=1e-5)
    assert torch.equal(bbox.cpu()[:, 6:], f[:, 7:9])
    assert torch.equal(prune.cpu()[:, 0], f[:, 1 + n_reg:].max(1).values)
    want = (torch.sigmoid(f[:, 1 + n_reg:]) * torch.sigmoid(f[:, :1])).max(1).values
    assert torch.allclose(mx.cpu(), want, atol=1e-6)


@pytest.mark.parametrize("n_reg,n_cls,n", [(6, 18, 70001), (8, 10, 3000), (6, 5, 129), (8, 23, 5)])
def test_fused_head_equals_gemm_then_head_post(dev, n_reg, n_cls, n):
    """b2_head_fused (head GEMM whose epilogue is the head post-processing) against the two-launch path b2_conv_tc ->
    b2_head_post on the same bf16 rows and weights: every output bit-identical (same accumulator, same expressions)."""
    g = torch.Generator().manual_seed(n)
    cols = 1 + n_reg + n_cls
    x = (torch.randn(n, 128, generator=g) * 0.7).to(dev).to(torch.bfloat16)
    w = torch.nn.functional.pad(torch.randn(128, cols, generator=g) * 0.05, (0, 32 - cols)).unsqueeze(0).to(dev).contiguous()
    b = torch.nn.functional.pad(torch.randn(cols, generator=g) * 0.1, (0, 32 - cols)).to(dev).contiguous()
    wt = w.transpose(1, 2).contiguous().to(torch.bfloat16)
    scale = torch.tensor([1.07], device=dev)
    assert ops.head_fused_ok(x, wt, n_reg, n_cls)
    got = ops.head_fused(x, wt, b, n_reg, n_cls, scale, want_max_score=True)
    h = ops.sparse_conv(x, w, None, n, wt_bf16=wt, bias=b, out_dtype=torch.float32)
    want = ops.head_post(h, n_reg, n_cls, scale, want_max_score=True)
    for a, e, name in zip(got, want, ("centerness", "bbox_pred", "cls", "prune", "max_score")):
        assert a.shape == e.shape and torch.equal(a, e), name
    # the persistent kernel (taken by size for >= 4 x SMs row tiles: the 70001-row case has 547) gives the same bits
    from fcaf3d_b200._lib import lib
    for mode in (1, 2):
        old = lib().b2_tc_persist_mode(mode)
        try:
            again = ops.head_fused(x, wt, b, n_reg, n_cls, scale, want_max_score=True)
        finally:
            lib().b2_tc_persist_mode(old)
        for a, e, name in zip(again, want, ("centerness", "bbox_pred", "cls", "prune", "max_score")):
            assert torch.equal(a, e), (mode, name)
