"""CPU: pins the oracle — against torch dense ops (self-checks that need no reference), against the reference's own
compiled CPU IoU (oracle/_ref), against known answers and against the committed golden vectors."""
import os

import numpy as np
import pytest
import torch

import oracle.me_cpu as OME
from oracle import nms_cpu, ref_loader
from conftest import GOLDEN


def _dense_grid(G, C, seed=0):
    torch.manual_seed(seed)
    g = torch.stack(torch.meshgrid(*[torch.arange(G)] * 3, indexing="ij"), -1).reshape(-1, 3)
    coords = torch.cat([torch.zeros(len(g), 1, dtype=torch.long), g], 1).int()
    return coords, torch.randn(len(g), C)


def test_conv_equals_conv3d_on_dense_grid():
    G = 6
    coords, f = _dense_grid(G, 4)
    x = OME.SparseTensor(f, coordinates=coords)
    conv = OME.MinkowskiConvolution(4, 5, kernel_size=3, dimension=3)
    y = conv(x).F
    w = conv.kernel.reshape(3, 3, 3, 4, 5).permute(4, 3, 2, 1, 0)  # k = ix + 3 iy + 9 iz -> [co, ci, ix, iy, iz]
    ref = torch.nn.functional.conv3d(f.reshape(G, G, G, 4).permute(3, 0, 1, 2)[None], w, padding=1)
    assert torch.allclose(y, ref[0].permute(1, 2, 3, 0).reshape(-1, 5), atol=1e-5)


def test_strided_conv_and_pool_on_dense_grid():
    G = 8
    coords, f = _dense_grid(G, 3, 1)
    x = OME.SparseTensor(f, coordinates=coords)
    pool = OME.MinkowskiMaxPooling(kernel_size=2, stride=2, dimension=3)
    y = pool(x)
    ref = torch.nn.functional.max_pool3d(f.reshape(G, G, G, 3).permute(3, 0, 1, 2)[None], 2)[0].permute(1, 2, 3, 0)
    c = y.C.long()
    assert y.tensor_stride == [2, 2, 2]
    assert torch.allclose(y.F, ref[c[:, 1] // 2, c[:, 2] // 2, c[:, 3] // 2], atol=0)
    # stride-2 3^3 conv == conv3d(stride 2) on the even sites (zero padding = absent neighbours)
    conv = OME.MinkowskiConvolution(3, 4, kernel_size=3, stride=2, dimension=3)
    z = conv(x)
    w = conv.kernel.reshape(3, 3, 3, 3, 4).permute(4, 3, 2, 1, 0)
    dense = torch.nn.functional.conv3d(f.reshape(G, G, G, 3).permute(3, 0, 1, 2)[None], w, padding=1)[0]
    c = z.C.long()
    assert torch.allclose(z.F, dense[:, c[:, 1], c[:, 2], c[:, 3]].t(), atol=1e-5)


def test_generative_transpose_equals_conv_transpose3d():
    G = 4
    coords, f = _dense_grid(G, 6, 2)
    coords[:, 1:] *= 2
    x = OME.SparseTensor(f, coordinates=coords, tensor_stride=2)
    up = OME.MinkowskiGenerativeConvolutionTranspose(6, 3, kernel_size=2, stride=2, dimension=3)
    y = up(x)
    assert len(y) == 8 * len(f) and y.tensor_stride == [1, 1, 1]
    w = up.kernel.reshape(2, 2, 2, 6, 3).permute(3, 4, 2, 1, 0)  # [ci, co, ix, iy, iz]
    ref = torch.nn.functional.conv_transpose3d(f.reshape(G, G, G, 6).permute(3, 0, 1, 2)[None], w, stride=2)[0]
    c = y.C.long()
    assert torch.allclose(y.F, ref[:, c[:, 1], c[:, 2], c[:, 3]].t(), atol=1e-5)


def test_dedup_first_occurrence_and_union_order():
    c = torch.tensor([[0, 5, 5, 5], [0, 1, 1, 1], [0, 5, 5, 5], [1, 1, 1, 1], [1, 1, 1, 1], [1, -3, 0, 2]]).int()
    f = torch.arange(6).float()[:, None]
    x = OME.SparseTensor(f, coordinates=c)
    assert x.unique_index.tolist() == [0, 1, 3, 5]
    assert x.inverse_mapping.tolist() == [0, 1, 0, 2, 2, 3]
    assert x.F[:, 0].tolist() == [0, 1, 3, 5]
    cm = x.coordinate_manager
    other = torch.tensor([[0, 1, 1, 1], [0, 9, 9, 9], [1, -3, 0, 2], [1, 7, 7, 7]]).int()
    key_b, _ = cm.insert_and_map(other, 1, "b")
    ukey, dst_a, dst_b = cm.union(x.coordinate_map_key, key_b)
    u = cm.get_map(ukey).coords
    assert u.tolist() == [[0, 5, 5, 5], [0, 1, 1, 1], [0, 9, 9, 9], [1, 1, 1, 1], [1, -3, 0, 2], [1, 7, 7, 7]]
    assert dst_a.tolist() == [0, 1, 3, 4] and dst_b.tolist() == [1, 2, 4, 5]


def test_interpolation_weights():
    c = torch.tensor([[0, 0, 0, 0], [0, 2, 0, 0]]).int()
    x = OME.SparseTensor(torch.tensor([[1.0], [3.0]]), coordinates=c, tensor_stride=2)
    q = torch.tensor([[0, 0, 0, 0], [0, 1, 0, 0], [0, 2, 0, 0], [0, 3, 0, 0], [0, 1, 1, 0.]])
    assert x.features_at_coordinates(q)[:, 0].tolist() == [1.0, 2.0, 3.0, 1.5, 1.0]


def test_nms_restatement_equals_reference_cpu_source(oracle_built):
    from fcaf3d_b200.scenes import random_boxes
    if "ref_cpu" not in oracle_built:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    for seed in range(3):
        b, _ = random_boxes(200, seed, rotated=True)
        ours = nms_cpu.pair_matrix(b, b, 1)
        ref = nms_cpu.ref_iou_bev_cpu(b, b)
        assert np.array_equal(np.nan_to_num(ours, nan=-1), np.nan_to_num(ref, nan=-1))


def test_nms_known_answers(oracle_built):
    x = np.array([[0, 0, 0, 2, 2, 1, 0], [1, 1, 0, 2, 2, 1, 0], [0, 0, 0, 2, 2, 1, np.pi / 2]], np.float32)
    iou = nms_cpu.pair_matrix(x, x, 1)
    assert abs(iou[0, 1] - 1 / 7) < 1e-6 and abs(iou[0, 2] - 1.0) < 1e-5
    assert abs(nms_cpu.pair_matrix(x, x, 2)[0, 1] - 1 / 7) < 1e-6
    # greedy loop vs brute force on axis-aligned boxes
    from fcaf3d_b200.scenes import random_boxes
    b, s = random_boxes(150, 3, rotated=False)
    b = b[np.argsort(-s)]
    keep = nms_cpu.nms_cpu(b, 0.3, False).tolist()
    m = nms_cpu.pair_matrix(b, b, 2)
    alive, brute = np.ones(len(b), bool), []
    for i in range(len(b)):
        if alive[i]:
            brute.append(i)
            alive &= ~(m[i] > 0.3) | (np.arange(len(b)) <= i)
    assert keep == brute
    assert nms_cpu.nms_cpu(b[:0], 0.5, True).tolist() == []
    assert nms_cpu.nms_cpu(b[:1], 0.5, True).tolist() == [0]


def test_axis_aligned_overlaps_docstring_golden():
    g = np.load(os.path.join(GOLDEN, "ops_ref.npz"))
    want = np.array([[0.25, 0.0, 0.0], [0.0, 0.0, 1.0], [0.0, 0.0, 0.0]], np.float32)
    assert np.allclose(g["aa_iou"][0, 0], 0.25) and np.allclose(g["aa_iou"][1, 2], 1.0)
    assert np.allclose(g["aa_iou"] > 0, want > 0)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree absent (GPU box)")
def test_golden_fixtures_match_a_fresh_reference_run():
    """The committed fixtures are what tests/golden/make_golden.py produces from the reference files today."""
    ns = ref_loader.load()
    g = np.load(os.path.join(GOLDEN, "ops_ref.npz"))
    pred, tgt, w = (torch.from_numpy(g[k]) for k in ("iou_pred", "iou_tgt", "iou_w"))
    loss = ns.IoU3DLoss(with_yaw=False, loss_weight=1.0)(pred[:, :6], tgt[:, :6], weight=w, avg_factor=w.sum())
    assert abs(loss.item() - float(g["iou_aa_loss"])) < 1e-6
    iou = ns.cal_iou_3d(pred[None], tgt[None])[0]
    assert torch.allclose(iou, torch.from_numpy(g["iou_rot_values"]), atol=1e-6)
    # rotated IoU of the reference's torch code agrees with the pcdet BEV overlap x height overlap
    bev = nms_cpu.pair_matrix(g["iou_pred"], g["iou_tgt"], 0).diagonal()
    p, t = g["iou_pred"], g["iou_tgt"]
    zo = np.clip(np.minimum(p[:, 2] + p[:, 5] / 2, t[:, 2] + t[:, 5] / 2) -
                 np.maximum(p[:, 2] - p[:, 5] / 2, t[:, 2] - t[:, 5] / 2), 0, None)
    inter = bev * zo
    iou3d = inter / (p[:, 3] * p[:, 4] * p[:, 5] + t[:, 3] * t[:, 4] * t[:, 5] - inter)
    # (pcdet counts corners within a 1 cm MARGIN as inside, the torch code does not: a few pairs differ by ~1e-2)
    diff = np.abs(iou3d - g["iou_rot_values"])
    assert np.quantile(diff, 0.8) < 1e-4 and diff.max() < 3e-2


def test_augmentation_oracle_matches_reference_containers():
    """oracle/data_cpu.py against vectors produced by the reference's own DepthInstance3DBoxes / DepthPoints methods
    (tests/golden/augment_ref.npz, generated by make_golden.py --augment-only); fp32 rounding only: 1e-6."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from oracle import data_cpu as D
    g = np.load(os.path.join(GOLDEN, "augment_ref.npz"))
    for ci in range(4):
        wy, fh, fv, ang, sc, tx, ty, tz = g["c%d_params" % ci]
        p, b = D.random_flip(g["c%d_points_in" % ci], g["c%d_boxes_in" % ci], bool(fh), bool(fv), bool(wy))
        p, b = D.global_rot_scale_trans(p, b, ang, sc, (tx, ty, tz), bool(wy))
        assert np.abs(p - g["c%d_points_out" % ci]).max() < 1e-6 * 4
        assert np.abs(b - g["c%d_boxes_out" % ci]).max() < 1e-6 * 4
    # sampling: without replacement when the scene is large enough, with replacement otherwise; reproducible draw order
    rng = np.random.default_rng(0)
    pts = np.arange(50, dtype=np.float32).reshape(50, 1).repeat(6, 1)
    s, ch = D.indoor_point_sample(pts, 20, rng)
    assert len(np.unique(ch)) == 20 and np.array_equal(s, pts[ch])
    s, ch = D.indoor_point_sample(pts, 80, rng)
    assert len(ch) == 80 and ch.max() < 50
    a, b = D.draw_params(np.random.default_rng(3), 1000, 100), D.draw_params(np.random.default_rng(3), 1000, 100)
    assert a["angle"] == b["angle"] and np.array_equal(a["choices"], b["choices"]) and abs(a["angle"]) <= 0.087266
