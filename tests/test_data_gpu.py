"""GPU: the device-side data path (fcaf3d_b200/data.py, b2_augment_points / b2_augment_boxes) against the numpy oracle
of the reference's transforms (oracle/data_cpu.py) and the golden vectors made with the reference's own containers.
fp32 tolerance 2e-6 absolute on coordinates of a few metres (one rounding per operation, FMA contraction may differ)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from fcaf3d_b200.data import DeviceAugment, ScenePrefetcher
from oracle import data_cpu as D

pytestmark = pytest.mark.gpu


def test_augment_kernels_match_reference_golden(dev):
    g = np.load(os.path.join(GOLDEN, "augment_ref.npz"))
    for ci in range(4):
        wy, fh, fv, ang, sc, tx, ty, tz = g["c%d_params" % ci]
        aug = DeviceAugment(num_points=0, with_yaw=bool(wy))
        params = dict(choices=None, flip_h=bool(fh), flip_v=bool(fv), angle=float(ang), scale=float(sc), trans=[tx, ty, tz])
        p, b, err = aug.apply(torch.from_numpy(g["c%d_points_in" % ci]).to(dev),
                              torch.from_numpy(g["c%d_boxes_in" % ci]).to(dev), params)
        assert int(err) == 0
        assert np.abs(p.cpu().numpy() - g["c%d_points_out" % ci]).max() < 2e-6
        assert np.abs(b.cpu().numpy() - g["c%d_boxes_out" % ci]).max() < 2e-6


@pytest.mark.parametrize("n_in,num_points,with_yaw", [(150000, 100000, False), (30000, 100000, True)])
def test_device_pipeline_matches_oracle_with_host_drawn_numbers(dev, n_in, num_points, with_yaw):
    rng = np.random.default_rng(5)
    pts = rng.uniform(-4, 4, (n_in, 6)).astype(np.float32)
    gt = np.concatenate([rng.uniform(-3, 3, (12, 3)), rng.uniform(0.3, 2, (12, 3)), rng.uniform(-1, 1, (12, 1))], 1).astype(np.float32)
    aug = DeviceAugment(num_points=num_points, with_yaw=with_yaw, seed=9, device_sampling=False)
    ref = D.draw_params(np.random.default_rng(9), n_in, num_points)
    params = aug.draw(n_in)
    assert np.array_equal(params["choices"], ref["choices"]) and params["angle"] == ref["angle"]      # same draw order
    p, b, err = aug.apply(torch.from_numpy(pts).to(dev), torch.from_numpy(gt).to(dev), params)
    want_p, want_b = D.random_flip(pts[ref["choices"]], gt, ref["flip_h"], ref["flip_v"], with_yaw)
    want_p, want_b = D.global_rot_scale_trans(want_p, want_b, ref["angle"], ref["scale"], ref["trans"], with_yaw)
    assert int(err) == 0 and p.shape == (num_points, 6)
    assert np.abs(p.cpu().numpy() - want_p).max() < 2e-6 and np.abs(b.cpu().numpy() - want_b).max() < 2e-6
    # colours ride along untouched
    assert np.array_equal(p.cpu().numpy()[:, 3:], pts[ref["choices"]][:, 3:])


def test_device_sampling_and_errors(dev):
    pts = torch.arange(5000, dtype=torch.float32, device=dev).reshape(-1, 1).repeat(1, 6)
    aug = DeviceAugment(num_points=2000, seed=1)                  # indices drawn on the device
    outs, boxes = aug([pts, pts[:1500]], [torch.zeros(0, 7, device=dev), torch.zeros(3, 7, device=dev)])
    assert outs[0].shape == (2000, 6) and outs[1].shape == (2000, 6) and boxes[0].shape == (0, 7)
    assert len(torch.unique(outs[0][:, 3])) == 2000               # without replacement when the scene is large enough
    assert outs[1][:, 3].max() < 1500                             # with replacement otherwise
    bad = dict(choices=torch.tensor([0, 7000], device=dev), flip_h=False, flip_v=False, angle=0.0, scale=1.0, trans=[0, 0, 0])
    _, _, err = aug.apply(pts, None, bad)
    assert int(err) == 1
    with pytest.raises(Exception):
        aug.apply(pts.cpu(), None, bad)                           # no CPU path


def test_scene_prefetcher_overlapped_staging(dev):
    rng = np.random.default_rng(0)
    batches = [[rng.normal(size=(1000 + 10 * i + j, 6)).astype(np.float32) for j in range(3)] for i in range(5)]
    got = []
    for batch in ScenePrefetcher(batches, dev):
        got.append([t.clone() for t in batch])
        torch.cuda.current_stream().synchronize()
    assert len(got) == 5
    for gb, hb in zip(got, batches):
        for t, h in zip(gb, hb):
            assert t.is_cuda and np.array_equal(t.cpu().numpy(), h)


@pytest.mark.parametrize("case", ["aa", "rot"])
def test_indoor_eval_matches_reference_golden(dev, case):
    """fcaf3d_b200.evaluation.indoor_eval (GPU IoU) against the reference's own indoor_eval.py
    (tests/golden/indoor_eval_ref.npz): every key of the result dict, 1e-5 absolute (nan where the reference has nan)."""
    from fcaf3d_b200.evaluation import indoor_eval
    g = np.load(os.path.join(GOLDEN, "indoor_eval_ref.npz"))
    n_img = len([k for k in g.files if k.startswith(case + "_gt") and not k.startswith(case + "_gtl")])
    gt_annos, dt_annos = [], []
    for i in range(n_img):
        gt, gl = g["%s_gt%d" % (case, i)], g["%s_gtl%d" % (case, i)]
        gt_annos.append(dict(gt_num=len(gt), gt_boxes_upright_depth=gt if case == "rot" else gt[:, :6], **{"class": gl}))
        dt_annos.append(dict(boxes_3d=torch.from_numpy(g["%s_det%d" % (case, i)]).to(dev),
                             scores_3d=torch.from_numpy(g["%s_dets%d" % (case, i)]),
                             labels_3d=torch.from_numpy(g["%s_detl%d" % (case, i)])))
    ret = indoor_eval(gt_annos, dt_annos, (0.25, 0.5), {c: "c%d" % c for c in range(5)})
    want = dict(zip([str(k) for k in g["%s_keys" % case]], g["%s_vals" % case]))
    assert set(ret) == set(want)
    for k, v in want.items():
        if np.isnan(v):
            assert np.isnan(ret[k]), k
        else:
            assert abs(ret[k] - v) < 1e-5, (k, ret[k], v)
