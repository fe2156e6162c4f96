"""CPU: host-side logic that needs no GPU — scene generator, name-keyed weights, state-dict compatibility of the
detector mirror with the reference's own model file loaded on our ME surface, config values, scene sharding."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_loader


def test_scene_generator_is_deterministic_and_scannet_shaped():
    from fcaf3d_b200.scenes import make_batch, make_scene
    p, gt, lab = make_scene("S0", 0)
    q, _, _ = make_scene("S0", 0)
    assert p.shape == (100000, 6) and p.dtype == np.float32 and np.array_equal(p, q)
    assert gt.shape == (15, 7) and lab.max() < 18
    vox = np.unique(np.floor(p[:, :3] / 0.01).astype(np.int64), axis=0)
    assert 95000 < len(vox) < 100000                       # ~1.7 % duplicate voxels at 1 cm
    assert p[:, 3:].min() >= 0 and p[:, 3:].max() <= 255
    batch = make_batch("tiny", 3, seed=1)
    assert len(batch) == 3 and not np.array_equal(batch[0][0], batch[1][0])
    s3, gt3, _ = make_scene("S3", 2)
    assert s3.shape[0] == 100000 and np.abs(gt3[:, 6]).max() > 0


def test_deterministic_init_is_order_independent():
    from fcaf3d_b200.detector import build_fcaf3d
    from fcaf3d_b200.testing import deterministic_init
    a = deterministic_init(build_fcaf3d("s3dis", depth=14), 3)
    torch.manual_seed(123)
    b = deterministic_init(build_fcaf3d("s3dis", depth=14), 3)
    for (ka, va), (kb, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)


def test_mirror_has_reference_hyperparameters():
    from fcaf3d_b200.detector import build_fcaf3d
    d = build_fcaf3d("scannet")
    h = d.neck_with_head
    assert (h.n_classes, h.n_reg_outs, h.pts_threshold, d.voxel_size) == (18, 6, 100000, 0.01)
    assert h.test_cfg == dict(nms_pre=1000, iou_thr=.5, score_thr=.01)
    assert abs(float(h.cls_conv.bias[0, 0]) + 4.59512) < 1e-4          # bias_init_with_prob(.01)
    assert sum(p.numel() for p in d.parameters()) == 70479574
    s = build_fcaf3d("sunrgbd").neck_with_head
    assert (s.n_classes, s.n_reg_outs, s.yaw_parametrization) == (10, 8, "fcaf3d")


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree absent (GPU box)")
def test_state_dict_and_config_match_the_reference_files():
    """Keys / shapes of the mirror == keys / shapes of the reference's own detector (built from its own files and
    its own config), so released checkpoints and the golden weights map one to one."""
    from fcaf3d_b200.detector import build_fcaf3d
    for cfg, ds in (("fcaf3d_scannet-3d-18class.py", "scannet"), ("fcaf3d_sunrgbd-3d-10class.py", "sunrgbd"),
                    ("fcaf3d_s3dis-3d-5class.py", "s3dis")):
        ref = ref_loader.build_reference_detector(cfg)
        ours = build_fcaf3d(ds)
        rs, os_ = ref.state_dict(), ours.state_dict()
        assert list(rs.keys()) == list(os_.keys())
        assert all(rs[k].shape == os_[k].shape for k in rs)
        c = ref_loader.load_config(cfg)["model"]
        assert c["neck_with_head"]["n_classes"] == ours.neck_with_head.n_classes
        assert c["test_cfg"] == ours.neck_with_head.test_cfg and c["voxel_size"] == ours.voxel_size


def _reference_root():
    from fcaf3d_b200 import compat
    from oracle import build_oracle
    for root in (ref_loader.REF, build_oracle.STAGED_PY):
        if compat.reference_files_present(root):
            return root
    return None


@pytest.mark.skipif(_reference_root() is None, reason="reference model files neither checked out nor staged")
def test_reference_model_files_build_on_the_product_surfaces(lib_built):
    """Drop-in check that needs no GPU: the reference's three model files, imported unmodified by fcaf3d_b200.compat
    (product ME surface, pcdet_nms, loss registry), construct the same module tree as the product's mirror, with the
    product's loss / NMS callables bound where the reference looks them up (a forward needs a GPU)."""
    import sys
    from fcaf3d_b200 import compat, losses, me, pcdet_nms
    from fcaf3d_b200.detector import build_fcaf3d
    saved = dict(sys.modules)
    try:
        ns = compat.load_reference_models(_reference_root())
        for dataset in ("scannet", "sunrgbd", "s3dis"):
            ref = compat.reference_detector(ns, dataset)
            ours = build_fcaf3d(dataset)
            rs, os_ = ref.state_dict(), ours.state_dict()
            assert list(rs.keys()) == list(os_.keys())
            assert all(rs[k].shape == os_[k].shape for k in rs)
            ref.load_state_dict(os_, strict=True)
            head = ref.neck_with_head
            assert isinstance(head.loss_cls, losses.FocalLoss) and isinstance(head.loss_bbox, losses.IoU3DLoss)
            assert head.loss_bbox.with_yaw == (dataset == "sunrgbd")
            assert type(head.assigner).__module__.endswith("fcaf3d_neck_with_head")     # the reference's own assigner
        net = ref.backbone
        assert isinstance(net.conv1[0], me.MinkowskiConvolution) and net.conv1[0].kernel.shape == (27, 3, 64)
        assert net.layer1[0].downsample[0].kernel.shape == (1, 64, 64)      # K == 1 but stride 2 -> [1, in, out]
        mod = sys.modules["mmdet3d.models.dense_heads.fcaf3d_neck_with_head"]
        assert mod.pcdet_nms_gpu is pcdet_nms.pcdet_nms_gpu and mod.ME is me
        assert open(mod.__file__, "rb").read() == open(
            os.path.join(_reference_root(), "mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py"), "rb").read()
    finally:
        for k in list(sys.modules):
            if k not in saved:
                del sys.modules[k]
        sys.modules.update(saved)


def test_checkpoint_round_trip_in_the_reference_format(tmp_path):
    """mmcv-style checkpoint (meta + state_dict, `module.` prefix of a DataParallel wrapper) -> strict load into a fresh
    mirror; a tensor of the wrong shape or a missing key is refused."""
    import collections
    import pytest
    import torch
    from fcaf3d_b200.checkpoint import load_checkpoint, save_checkpoint
    from fcaf3d_b200.detector import build_fcaf3d
    from fcaf3d_b200.testing import deterministic_init
    src = deterministic_init(build_fcaf3d("s3dis", depth=14), 5)
    path = str(tmp_path / "epoch_12.pth")
    save_checkpoint(src, path, meta=dict(epoch=12, CLASSES=("table", "chair")))
    dst = build_fcaf3d("s3dis", depth=14)
    meta = load_checkpoint(dst, path)
    assert meta["epoch"] == 12
    for (ka, va), (kb, vb) in zip(src.state_dict().items(), dst.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)
    wrapped = {"meta": {}, "state_dict": collections.OrderedDict(("module." + k, v) for k, v in src.state_dict().items())}
    torch.save(wrapped, path)
    dst2 = build_fcaf3d("s3dis", depth=14)
    load_checkpoint(dst2, path)
    assert torch.equal(dst2.state_dict()["backbone.conv1.0.kernel"], src.state_dict()["backbone.conv1.0.kernel"])
    sd = dict(src.state_dict())
    sd["backbone.conv1.0.kernel"] = sd["backbone.conv1.0.kernel"][:, :, :32]
    torch.save({"state_dict": sd}, path)
    with pytest.raises(RuntimeError, match="another shape"):
        load_checkpoint(build_fcaf3d("s3dis", depth=14), path)
    sd = dict(src.state_dict())
    sd.pop("neck_with_head.cls_conv.bias")
    torch.save({"state_dict": sd}, path)
    with pytest.raises(RuntimeError, match="missing"):
        load_checkpoint(build_fcaf3d("s3dis", depth=14), path)
