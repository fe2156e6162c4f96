"""GPU: ME-surface completeness and edge cases beyond the default FCAF3D config — TensorField, 2-/3-scale variants,
Bottleneck backbones, batches with an empty or tiny scene, negative / far coordinates."""
import numpy as np
import pytest
import torch

import oracle.me_cpu as OME
from fcaf3d_b200 import me as ME
from fcaf3d_b200.detector import MEResNet3D, build_fcaf3d
from fcaf3d_b200.scenes import make_scene
from fcaf3d_b200.testing import deterministic_init
from util import reference_detections, reference_twin, rel_err

pytestmark = pytest.mark.gpu


def test_tensor_field_quantise_and_slice(dev):
    g = torch.Generator().manual_seed(0)
    xyz = torch.rand(5000, 3, generator=g) * 20 - 10
    coords = torch.cat([torch.zeros(5000, 1), xyz], 1).to(dev)
    feats = torch.randn(5000, 4, generator=g).to(dev)
    tf = ME.TensorField(feats, coords)
    st = tf.sparse()
    # UNWEIGHTED_AVERAGE: voxel feature = mean of its points
    inv = tf.inverse_mapping
    want = torch.zeros(len(st), 4, device=dev).index_add_(0, inv, feats)
    want /= torch.bincount(inv, minlength=len(st)).unsqueeze(1)
    assert torch.allclose(st.F, want, atol=1e-5)
    assert torch.equal(st.C[inv], torch.floor(coords).int())
    back = st.slice(tf)
    assert back.F.shape == feats.shape and torch.equal(back.F, st.F[inv])


@pytest.mark.parametrize("n_outs,voxel", [(2, 0.02), (3, 0.01)])
def test_two_and_three_scale_variants(dev, ref_models, n_outs, voxel):
    """configs/fcaf3d/fcaf3d_{2,3}scales_scannet-3d-18class.py: fewer levels (and 2 cm voxels for the 2-scale one)"""
    det = deterministic_init(build_fcaf3d("scannet", n_outs=n_outs, voxel_size=voxel), 0, cls_bias=0.0).to(dev).eval()
    pts = [torch.from_numpy(make_scene("tiny", s, 3000)[0]).to(dev) for s in (1, 2)]
    ME.set_compute_dtype(torch.float32)
    with torch.no_grad():
        fused = det.simple_test_fused(pts)
        twin = reference_twin(ref_models, det, "scannet", n_outs=n_outs, voxel_size=voxel)
        ref = reference_detections(twin, [list(a) for a in twin.extract_feat(pts, None)], 2)
    assert len(fused) == 2
    for (b, s, l), (rb, rs, rl) in zip(fused, ref):
        assert b.shape[1] == 6 and abs(len(b) - len(rb)) <= 2 and len(b) > 0


def test_bottleneck_backbone_runs_and_matches_unfused(dev):
    net = MEResNet3D(3, 50, n_outs=2).to(dev).eval()
    net.init_weights()
    pts = make_scene("tiny", 3, 2000)[0]
    c = torch.from_numpy(np.concatenate([np.zeros((len(pts), 1)), np.floor(pts[:, :3] / 0.02)], 1)).int().to(dev)
    x = ME.SparseTensor(torch.from_numpy(pts[:, 3:] / 255).float().to(dev), coordinates=c)
    with torch.no_grad():
        a = net(x)
        b = net.forward_fused(ME.SparseTensor(x.F, coordinate_map_key=x.coordinate_map_key,
                                              coordinate_manager=x.coordinate_manager))
    assert [t.F.shape[1] for t in a] == [256, 512]
    for u, v in zip(a, b):
        assert torch.equal(u.C, v.C) and rel_err(u.F, v.F) < 1e-4


def test_batch_with_an_empty_and_a_tiny_scene(dev):
    det = deterministic_init(build_fcaf3d("scannet", depth=14), 0, cls_bias=0.0).to(dev).eval()
    full = torch.from_numpy(make_scene("tiny", 4, 3000)[0]).to(dev)
    tiny = full[:40].contiguous()
    empty = full[:0].contiguous()
    ME.set_compute_dtype(torch.float32)
    with torch.no_grad():
        out = det.simple_test_fused([full, empty, tiny])
        alone = det.simple_test_fused([full])
    assert len(out) == 3 and len(out[1][0]) == 0
    # scenes do not interact (InstanceNorm is per scene, BatchNorm uses running statistics in eval mode)
    assert len(out[0][0]) == len(alone[0][0])
    key = lambda t: sorted(zip(t[2].tolist(), [round(float(s), 5) for s in t[1].tolist()]))
    assert key(out[0]) == key(alone[0])


def test_negative_and_far_coordinates_match_oracle(dev):
    g = torch.Generator().manual_seed(5)
    c = torch.randint(-60000, 60000, (4000, 3), generator=g)
    c[:2000] = c[:2000] // 4000 * 4  # a dense cluster with real neighbourhoods around the origin
    coords = torch.cat([torch.zeros(4000, 1, dtype=torch.long), c], 1).int()
    f = torch.randn(4000, 8, generator=g)
    o = OME.SparseTensor(f, coordinates=coords)
    x = ME.SparseTensor(f.to(dev), coordinates=coords.to(dev))
    oc = OME.MinkowskiConvolution(8, 16, kernel_size=3, stride=2, dimension=3)
    gc = ME.MinkowskiConvolution(8, 16, kernel_size=3, stride=2, dimension=3)
    gc.load_state_dict(oc.state_dict())
    with torch.no_grad():
        y, ry = gc.to(dev)(x), oc(o)
    assert torch.equal(y.C.cpu(), ry.C) and rel_err(y.F, ry.F) < 2e-5


def test_mefpn3d_matches_reference_golden(dev):
    """fcaf3d_b200.necks.MEFPN3D against the reference's own necks/me_fpn.py run on the oracle
    (tests/golden/mefpn_tiny.npz): evaluation outputs of the four levels (row sets exact, features 2e-3 of the column
    max) and one training forward/backward (pruning loss 2e-3, sampled gradients 1e-2 of max)."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from fcaf3d_b200.detector import MEResNet3D
    from fcaf3d_b200.necks import MEFPN3D
    from fcaf3d_b200.testing import canonical_order, deterministic_init
    g = np.load(os.path.join(GOLDEN, "mefpn_tiny.npz"))
    model = torch.nn.Module()
    model.backbone = MEResNet3D(in_channels=3, depth=34)
    model.neck = MEFPN3D(in_channels=(64, 128, 256, 512), out_channels=128, voxel_size=.01, pts_threshold=400)
    deterministic_init(model, seed=2)
    model = model.to(dev)
    nb = len(g["seeds"])
    scenes = [torch.from_numpy(g["points_%d" % i]).to(dev) for i in range(nb)]

    def feats():
        coords, f = ME.utils.batch_sparse_collate([(p[:, :3] / .01, p[:, 3:] / 255.) for p in scenes], device=dev)
        return model.backbone(ME.SparseTensor(coordinates=coords, features=f))

    ME.set_compute_dtype(torch.float32)
    model.eval()
    with torch.no_grad():
        outs = model.neck(feats(), img_metas=[dict()] * nb)
    for lvl, o in enumerate(outs):
        order = canonical_order(o.C.cpu())
        c, f = o.C.cpu()[order], o.F.float().cpu()[order]
        assert len(c) == int(g["eval_l%d_n" % lvl])
        step = max(1, len(c) // 256)
        assert np.array_equal(c[::step].numpy(), g["eval_l%d_coords" % lvl])
        want = torch.from_numpy(g["eval_l%d_feat" % lvl])
        scale = want.abs().max(0).values.clamp(min=1e-6)
        assert ((f[::step] - want).abs() / scale).max().item() < 2e-3
    model.train()
    gtb = [torch.from_numpy(g["gt_%d" % i]).to(dev) for i in range(nb)]          # gravity centre, size, yaw
    outs, losses = model.neck(feats(), gtb, None, [dict()] * nb)
    assert [len(o.F) for o in outs] == g["train_rows"].tolist()
    want_loss = float(g["loss_pruning"])
    assert abs(losses["loss_pruning"].item() - want_loss) < 2e-3 * abs(want_loss)
    (losses["loss_pruning"] + sum(o.F.square().mean() for o in outs)).backward()
    params = dict(model.named_parameters())
    for key in [k[5:] for k in g.files if k.startswith("grad/")]:
        flat = params[key].grad.reshape(-1)
        step = max(1, flat.numel() // 4096)
        want = torch.from_numpy(g["grad/" + key])
        err = ((flat[::step].cpu() - want).abs().max() / want.abs().max().clamp(min=1e-12)).item()
        assert err < 1e-2, (key, err)


@pytest.mark.parametrize("dtype,thr", [(torch.float32, 4000), (torch.bfloat16, 4000), (torch.float32, 700)])
def test_prune_before_up_block_convolution_gives_the_same_rows(dev, ref_models, dtype, thr):
    """forward_fused, eval: the up-block's 3^3 convolution evaluated at the surviving rows only (tables derived from the
    parent's) against convolve-everything-then-prune: same coordinates, bit-identical head outputs; the derived table of
    the pruned map equals the hash-built one; and the reference's own files, run op by op, agree within the fp32 tolerance.
    thr = 4000: only the finest level prunes (the shape of the real workloads); thr = 700: the ts16 level prunes too, after
    which backbone rows whose parents were dropped reach the union and the finest level takes the general path."""
    from fcaf3d_b200 import detector as D
    from fcaf3d_b200 import ops
    det = deterministic_init(build_fcaf3d("scannet", pts_threshold=thr), 0, cls_bias=0.0).to(dev).eval()
    assert det.neck_with_head.pts_threshold == thr
    pts = [torch.from_numpy(make_scene("tiny", s, 3000)[0]).to(dev) for s in (1, 2, 3)]
    ME.set_compute_dtype(dtype)
    outs = {}
    try:
        for flag in (True, False):
            D.PRUNE_BEFORE_CONV = flag
            with torch.no_grad():
                levels, cm = det.forward_features_fused(pts)
            outs[flag] = (levels, cm)
        with torch.no_grad():
            ref = (reference_twin(ref_models, det, "scannet", pts_threshold=thr).extract_feat(pts, None)
                   if dtype == torch.float32 else None)
    finally:
        D.PRUNE_BEFORE_CONV = True
        ME.set_compute_dtype(torch.float32)
    (la, cma), (lb, cmb) = outs[True], outs[False]
    pruned_levels = 0
    for a, b in zip(la, lb):
        ma, mb = cma.get_map(a["key"]), cmb.get_map(b["key"])
        assert torch.equal(ma.coords, mb.coords) and ma.off_host == mb.off_host
        for name in ("centerness", "bbox_pred", "cls", "prune"):
            assert torch.equal(a[name], b[name]), name
        ka = cma._kmaps.get((a["key"], a["key"], 3, 1))
        kb = cmb._kmaps.get((b["key"], b["key"], 3, 1))
        assert torch.equal(ka.nbr[:, :ka.n_out], kb.nbr[:, :kb.n_out])
        assert torch.equal(ka.tile_mask, kb.tile_mask)
        if max(ma.off_host[i + 1] - ma.off_host[i] for i in range(3)) == thr:
            pruned_levels += 1
            table, cap = ma.ensure_table()
            hashed = ops.kernel_map(table, cap, ma.coords, 3, ma.ts)
            assert torch.equal(hashed.nbr[:, :ka.n_out], ka.nbr[:, :ka.n_out])
            assert torch.equal(hashed.tile_mask, ka.tile_mask)
    assert pruned_levels >= 1 and det.neck_with_head.prune_first_levels >= 1
    if ref is not None:
        centernesses, bbox_preds, cls_scores, points = [list(x) for x in ref]
        for lvl, a in enumerate(la):
            m = cma.get_map(a["key"])
            for b in range(3):
                sl = slice(m.off_host[b], m.off_host[b + 1])
                assert torch.equal(torch.round(points[lvl][b] / det.voxel_size).int(), m.coords[sl, 1:])
                want = cls_scores[lvl][b]
                assert ((a["cls"][sl] - want).abs().max() / want.abs().max().clamp(min=1e-6)).item() < 2e-3
