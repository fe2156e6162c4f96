"""CPU: the C-ABI library loads and exports exactly what include/fcaf3d_b200.h declares; host-only entry points."""
import ctypes
import re

from fcaf3d_b200 import _lib


def test_header_parses_and_every_symbol_is_exported(lib_built):
    protos = _lib.parse_header()
    assert len(protos) >= 38
    raw = ctypes.CDLL(lib_built)
    for name in protos:
        assert hasattr(raw, name), "libfcaf3d_b200.so lacks %s declared in the header" % name


def test_no_undeclared_extern_c_entry_points(lib_built):
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", lib_built], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (b2_\w+)$", out, flags=re.M))
    declared = set(_lib.parse_header())
    assert exported == declared, (exported - declared, declared - exported)


def test_host_helpers_and_argument_errors(lib_built):
    l = _lib.lib()
    assert l.b2_version() >= 100
    assert l.b2_scan_ws_bytes(10000) >= 8 * 3
    assert l.b2_map_build_ws_bytes(1000) > 9000
    # argument validation happens before any CUDA call, so it is testable without a GPU
    assert l.b2_table_clear(None, 1000, None) == -1
    assert "power of two" in _lib.last_error()
    assert l.b2_kernel_map(None, 1024, None, 10, 3, 1, None, 100, None, None) == -1
    assert l.b2_conv_tc(None, 10, None, None, 0, None, None, 10, 48, 64, 1, None, None, None, None, 0, 0, None, 1, None) == -1
    assert "cin" in _lib.last_error()


def test_nms_plan_host():
    from fcaf3d_b200 import ops
    tile_off, mask_off, words, mx = ops.nms_plan([0, 0, 64, 65 + 64, 129 + 4000])
    # segments: 0, 64, 65, 4000 boxes -> T = 0, 1, 2, 63 ; tiles T(T+1)/2
    assert tile_off == [0, 0, 1, 4, 4 + 63 * 64 // 2]
    assert mask_off == [0, 0, 64, 64 + 130, 64 + 130 + 4000 * 63]
    assert words == mask_off[-1] and mx == 4000


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    import pytest
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.B2Error, match="no CPU or PyTorch fallback"):
        _lib.lib()


def test_cpu_tensors_are_rejected(lib_built):
    import pytest
    import torch
    from fcaf3d_b200 import me
    with pytest.raises(_lib.B2Error):
        me.SparseTensor(torch.zeros(4, 3), coordinates=torch.zeros(4, 4, dtype=torch.int32))
