"""CPU: the rotated-NMS per-pair arithmetic must be contracted into FMAs identically in every kernel that inlines it
(see tools/check_nms_codegen.py); guards the bit-exactness that the GPU tests establish for k_nms_mask<true>."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("nvdisasm") is None or shutil.which("nvcc") is None, reason="needs nvcc + nvdisasm")
def test_rotated_nms_codegen_is_identical_across_kernels():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check_nms_codegen.py")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
