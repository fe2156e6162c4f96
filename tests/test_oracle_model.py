"""CPU: the oracle's stand-alone model restatement (oracle/model_cpu.py, the CPU baseline of bench.py) reproduces
the golden vectors that the reference's own model files produced (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from fcaf3d_b200.testing import canonical_order, deterministic_init
from oracle import model_cpu


@pytest.mark.parametrize("name,dataset", [("fcaf3d_tiny_sunrgbd.npz", "sunrgbd")])
def test_model_restatement_equals_reference_golden(name, dataset):
    g = np.load(os.path.join(GOLDEN, name))
    det = model_cpu.Detector(dataset)
    deterministic_init(det, seed=0, cls_bias=0.0)
    det.eval()
    pts = [torch.from_numpy(g["points_%d" % i]) for i in range(len(g["seeds"]))]
    levels = det(pts)
    for lvl, l in enumerate(levels):
        for b in range(len(pts)):
            sel = l["coords"][:, 0] == b
            vox = l["coords"][sel][:, 1:]
            feat = torch.cat([l["centerness"], l["bbox_pred"], l["cls"]], 1)[sel]
            order = canonical_order(vox)
            vox, feat = vox[order], feat[order]
            assert len(vox) == int(g["l%d_b%d_n" % (lvl, b)])
            step = max(1, len(vox) // 256)
            assert np.array_equal(vox[::step].numpy(), g["l%d_b%d_vox" % (lvl, b)])
            assert torch.allclose(feat[::step], torch.from_numpy(g["l%d_b%d_feat" % (lvl, b)]), rtol=1e-4, atol=1e-5)
