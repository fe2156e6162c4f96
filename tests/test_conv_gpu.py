"""GPU parity (floating point): sparse convolution (CUDA-core fp32 path and tcgen05 bf16 path), generative
transpose, pooling, normalisation and activations vs the CPU oracle.

Tolerances (relative to the max |reference| of the tensor):
  fp32 path  : 2e-5  (fp32 FMA accumulation in a different order than the oracle's index_add)
  bf16 path  : 1.5e-2 (operands rounded to bf16 = 2^-8 relative, fp32 accumulate, bf16 output rounding)
"""
import numpy as np
import pytest
import torch

import oracle.me_cpu as OME
from fcaf3d_b200 import me as ME
from fcaf3d_b200 import ops
from util import rel_err, surface_coords

pytestmark = pytest.mark.gpu
FP32_TOL, BF16_TOL = 2e-5, 1.5e-2


def _tensors(dev, c_in, batch=2, n_points=3000, seed=0, ts=1):
    c = torch.from_numpy(surface_coords(n_points, batch, seed))
    o = OME.SparseTensor(torch.zeros(len(c), 1), coordinates=c)
    n = len(o)
    torch.manual_seed(seed)
    f = torch.randn(n, c_in)
    o = OME.SparseTensor(f, coordinate_map_key=o.coordinate_map_key, coordinate_manager=o.coordinate_manager)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c.to(dev))
    assert torch.equal(x0.C.cpu(), o.C)
    x = ME.SparseTensor(f.to(dev), coordinate_map_key=x0.coordinate_map_key, coordinate_manager=x0.coordinate_manager)
    return x, o


def _copy_params(dst, src):
    dst.load_state_dict(src.state_dict())
    return dst


@pytest.mark.parametrize("cin,cout,k,stride,bias", [(3, 64, 3, 2, False), (16, 24, 3, 1, True), (64, 64, 3, 1, False),
                                                     (32, 18, 1, 1, True), (64, 128, 1, 2, False), (5, 7, 2, 2, False)])
def test_conv_fp32_path(dev, cin, cout, k, stride, bias):
    x, o = _tensors(dev, cin)
    oc = OME.MinkowskiConvolution(cin, cout, kernel_size=k, stride=stride, bias=bias, dimension=3)
    gc = _copy_params(ME.MinkowskiConvolution(cin, cout, kernel_size=k, stride=stride, bias=bias, dimension=3), oc).to(dev)
    with torch.no_grad():
        y, ry = gc(x), oc(o)
    assert torch.equal(y.C.cpu(), ry.C)
    assert y.F.dtype == torch.float32 and rel_err(y.F, ry.F) < FP32_TOL


@pytest.mark.parametrize("cin,cout,k,stride", [(64, 64, 3, 1), (64, 128, 3, 2), (128, 128, 3, 1), (256, 64, 1, 2),
                                               (128, 256, 3, 1)])
def test_conv_tcgen05_path(dev, cin, cout, k, stride):
    x, o = _tensors(dev, cin, seed=1)
    oc = OME.MinkowskiConvolution(cin, cout, kernel_size=k, stride=stride, dimension=3)
    gc = _copy_params(ME.MinkowskiConvolution(cin, cout, kernel_size=k, stride=stride, dimension=3), oc).to(dev)
    xb = ME.SparseTensor(x.F.to(torch.bfloat16), coordinate_map_key=x.coordinate_map_key,
                         coordinate_manager=x.coordinate_manager)
    assert ops.tc_eligible(xb.F, cin, cout)
    with torch.no_grad():
        y, ry = gc(xb), oc(o)
    torch.cuda.synchronize()
    assert y.F.dtype == torch.bfloat16
    assert rel_err(y.F, ry.F) < BF16_TOL
    # same kernel, fp32 output and the fused epilogue (bias-free conv -> scale/shift -> residual -> ReLU)
    out_key, km = gc.maps(xb)
    n_out = x.coordinate_manager.size(out_key)
    torch.manual_seed(2)
    scale, shift = torch.rand(cout, device=dev) + .5, torch.randn(cout, device=dev)
    res = torch.randn(n_out, cout, device=dev)
    z = ops.sparse_conv(xb.F, gc._w3(), km, n_out, wt_bf16=gc._wt_bf16(), scale=scale, shift=shift, residual=res,
                        act=ops.ACT_RELU, out_dtype=torch.float32)
    ref = torch.relu(ry.F * scale.cpu() + shift.cpu() + res.cpu())
    assert rel_err(z, ref) < BF16_TOL
    # and against the CUDA-core kernel fed the same bf16-rounded operands: only accumulation order differs
    ops.set_tensor_cores(False)
    try:
        z_cc = ops.sparse_conv(xb.F, gc._w3().to(torch.bfloat16).float(), km, n_out, scale=scale, shift=shift,
                               residual=res, act=ops.ACT_RELU, out_dtype=torch.float32)
    finally:
        ops.set_tensor_cores(True)
    assert rel_err(z, z_cc) < 5e-4


@pytest.mark.parametrize("cin,cout,stride,out_dtype", [(64, 64, 1, torch.bfloat16), (128, 128, 1, torch.float32),
                                                        (64, 128, 2, torch.float32)])
def test_sparsity_sorted_walk_is_bit_identical(dev, cin, cout, stride, out_dtype, monkeypatch):
    """b2_kmap_sort: perm is a permutation of the output rows, the sorted table is the permuted table, its tile masks
    are exact and have fewer offsets than the unsorted ones; the convolution over it equals the unsorted walk bit for
    bit (same per-row accumulation order), with the fused epilogue and a residual."""
    from util import surface_coords
    c = torch.from_numpy(surface_coords(9000, 2, seed=3)).to(dev)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c)
    cm, key = x0.coordinate_manager, x0.coordinate_map_key
    okey = key if stride == 1 else cm.stride(key, stride)
    km = cm.kernel_map(key, okey, 3)
    n_in, n_out = cm.size(key), cm.size(okey)
    perm, nbr_s, tm_s = ops.kmap_sorted(km)
    p = perm[:n_out].long()
    assert torch.equal(torch.sort(p).values, torch.arange(n_out, device=dev)) and bool((perm[n_out:] == -1).all())
    assert torch.equal(nbr_s[:, :n_out], km.nbr[:, p]) and bool((nbr_s[:, n_out:] == -1).all())
    present = (nbr_s >= 0).reshape(27, -1, 128).any(2).long()
    want = (present << torch.arange(27, device=dev).reshape(27, 1)).sum(0)
    assert torch.equal(tm_s.long() & 0x7FFFFFF, want)
    popc = lambda t: sum(((t.long() >> b) & 1).sum().item() for b in range(27))   # noqa: E731
    assert popc(tm_s) < popc(km.tile_mask)
    torch.manual_seed(0)
    f = torch.randn(n_in, cin, device=dev).to(torch.bfloat16)
    w = torch.randn(27, cin, cout, device=dev) * 0.05
    wt = w.transpose(1, 2).contiguous().to(torch.bfloat16)
    scale, shift = torch.rand(cout, device=dev) + .5, torch.randn(cout, device=dev)
    res = torch.randn(n_out, cout, device=dev).to(out_dtype)
    kw = dict(wt_bf16=wt, scale=scale, shift=shift, residual=res, act=ops.ACT_RELU, out_dtype=out_dtype)
    monkeypatch.setattr(ops, "SORT_MIN_ROWS", 1)
    monkeypatch.setattr(ops, "SORT_MIN_USES", 1)
    y_sorted = ops.sparse_conv(f, w, km, n_out, **kw)
    km.sorted = None
    monkeypatch.setattr(ops, "SORT_MIN_ROWS", 0)
    y_plain = ops.sparse_conv(f, w, km, n_out, **kw)
    assert km.sorted is None and torch.equal(y_sorted, y_plain)


@pytest.mark.parametrize("cin,cout,stride,dense", [(64, 64, 1, False), (128, 128, 1, False), (64, 128, 2, False),
                                                   (128, 64, 1, False), (256, 128, 1, False), (512, 128, 2, False),
                                                   (64, 256, 1, True), (256, 128, 1, True)])
def test_persistent_kernel_is_bit_identical(dev, cin, cout, stride, dense, monkeypatch):
    """csrc/conv_tcp.cu (CTAs walk tiles, two TMEM accumulators, epilogue warps one tile behind) against k_conv_tc on
    the same layer: equal bit for bit, with and without the sparsity-sorted table, fused epilogue + residual, a row count
    that is not a multiple of 128 and more tiles than resident CTAs (so that CTAs take several tiles); the identity map
    (generative transposes) with column tiles and ch_mod."""
    from fcaf3d_b200._lib import lib
    from util import surface_coords
    c = torch.from_numpy(surface_coords(60000, 3, seed=5)).to(dev)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c)
    cm, key = x0.coordinate_manager, x0.coordinate_map_key
    okey = key if stride == 1 else cm.stride(key, stride)
    n_in, n_out = cm.size(key), cm.size(okey)
    assert n_out > 300 * 128 or stride > 1
    torch.manual_seed(1)
    f = torch.randn(n_in, cin, device=dev).to(torch.bfloat16)
    kvol = 1 if dense else 27
    w = torch.randn(kvol, cin, cout, device=dev) * 0.05
    wt = w.transpose(1, 2).contiguous().to(torch.bfloat16)
    ch = 64 if dense else cout
    scale, shift = torch.rand(ch, device=dev) + .5, torch.randn(ch, device=dev)
    res = torch.randn(n_out, cout, device=dev).to(torch.bfloat16)
    outs = {}
    for mode in (1, 2):
        old = lib().b2_tc_persist_mode(mode)
        try:
            if dense:
                outs[mode] = (ops.sparse_conv(f, w, None, n_in, wt_bf16=wt, scale=scale, shift=shift, act=ops.ACT_ELU,
                                              ch_mod=64, out_dtype=torch.bfloat16),)
            else:
                km = cm.kernel_map(key, okey, 3)
                km.sorted = None
                kw = dict(wt_bf16=wt, scale=scale, shift=shift, residual=res, act=ops.ACT_RELU, out_dtype=torch.bfloat16)
                monkeypatch.setattr(ops, "SORT_MIN_ROWS", 0)
                plain = ops.sparse_conv(f, w, km, n_out, **kw)
                monkeypatch.setattr(ops, "SORT_MIN_ROWS", 1)
                monkeypatch.setattr(ops, "SORT_MIN_USES", 1)
                outs[mode] = (plain, ops.sparse_conv(f, w, km, n_out, **kw))
                assert km.sorted is not None
        finally:
            lib().b2_tc_persist_mode(old)
    for a, b in zip(outs[1], outs[2]):
        assert torch.equal(a, b)
    if not dense:
        assert torch.equal(outs[2][0], outs[2][1])


def test_conv_tcgen05_wide_tiles(dev):
    """cout % 256 == 0 on a map with >= 74 row tiles takes the 256-column tile (one CTA per SM); same checks as above:
    against the CUDA-core kernel on the same bf16-rounded operands (5e-4) with the fused epilogue."""
    from util import surface_coords
    c = torch.from_numpy(surface_coords(16000, 2, seed=8)).to(dev)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c)
    cm, key = x0.coordinate_manager, x0.coordinate_map_key
    n = cm.size(key)
    assert (n + 127) // 128 >= 74
    km = cm.kernel_map(key, key, 3)
    torch.manual_seed(4)
    for cin, cout in ((64, 256), (128, 512)):
        f = torch.randn(n, cin, device=dev).to(torch.bfloat16)
        w = (torch.randn(27, cin, cout, device=dev) * 0.05).to(torch.bfloat16).float()
        wt = w.transpose(1, 2).contiguous().to(torch.bfloat16)
        scale, shift = torch.rand(cout, device=dev) + .5, torch.randn(cout, device=dev)
        res = torch.randn(n, cout, device=dev)
        kw = dict(scale=scale, shift=shift, residual=res, act=ops.ACT_ELU, out_dtype=torch.float32)
        z = ops.sparse_conv(f, w, km, n, wt_bf16=wt, **kw)
        ops.set_tensor_cores(False)
        try:
            z_cc = ops.sparse_conv(f, w, km, n, **kw)
        finally:
            ops.set_tensor_cores(True)
        assert rel_err(z, z_cc) < 5e-4
        zb = ops.sparse_conv(f, w, km, n, wt_bf16=wt, scale=scale, shift=shift, act=ops.ACT_RELU, out_dtype=torch.bfloat16)
        ref_b = torch.relu(ops.sparse_conv(f, w, km, n, wt_bf16=wt, scale=scale, shift=shift, out_dtype=torch.float32))
        assert rel_err(zb, ref_b) < 1e-2


def test_generative_transpose_both_paths(dev):
    x, o = _tensors(dev, 128, n_points=1500)
    cm, ocm = x.coordinate_manager, o.coordinate_manager
    k8, ok8 = cm.stride(cm.stride(cm.stride(x.coordinate_map_key, 2), 2), 2), None
    ok8 = ocm.stride(ocm.stride(ocm.stride(o.coordinate_map_key, 2), 2), 2)
    n = cm.size(k8)
    torch.manual_seed(3)
    f = torch.randn(n, 128)
    xs = ME.SparseTensor(f.to(dev), coordinate_map_key=k8, coordinate_manager=cm)
    os_ = OME.SparseTensor(f, coordinate_map_key=ok8, coordinate_manager=ocm)
    ou = OME.MinkowskiGenerativeConvolutionTranspose(128, 64, kernel_size=2, stride=2, dimension=3)
    gu = _copy_params(ME.MinkowskiGenerativeConvolutionTranspose(128, 64, kernel_size=2, stride=2, dimension=3), ou).to(dev)
    with torch.no_grad():
        y, ry = gu(xs), ou(os_)
        yb = gu(ME.SparseTensor(xs.F.to(torch.bfloat16), coordinate_map_key=k8, coordinate_manager=cm))
    assert torch.equal(y.C.cpu(), ry.C) and y.tensor_stride == [4, 4, 4]
    assert rel_err(y.F, ry.F) < FP32_TOL
    assert rel_err(yb.F, ry.F) < BF16_TOL


def test_pool_norms_activations(dev):
    x, o = _tensors(dev, 64, batch=3)
    with torch.no_grad():
        y, ry = ME.MinkowskiMaxPooling(2, 2, dimension=3)(x), OME.MinkowskiMaxPooling(2, 2, dimension=3)(o)
        assert torch.equal(y.C.cpu(), ry.C) and torch.equal(y.F.cpu(), ry.F)
        oin = OME.MinkowskiInstanceNorm(64)
        torch.nn.init.normal_(oin.weight), torch.nn.init.normal_(oin.bias)
        gin = _copy_params(ME.MinkowskiInstanceNorm(64), oin).to(dev)
        assert rel_err(gin(x).F, oin(o).F) < 1e-4
        obn = OME.MinkowskiBatchNorm(64)
        torch.nn.init.normal_(obn.bn.weight), torch.nn.init.normal_(obn.bn.bias)
        gbn = _copy_params(ME.MinkowskiBatchNorm(64), obn).to(dev)
        gbn.train(), obn.train()
        assert rel_err(gbn(x).F, obn(o).F) < 1e-4
        assert rel_err(gbn.bn.running_var, obn.bn.running_var) < 1e-5
        assert rel_err(gbn.bn.running_mean, obn.bn.running_mean) < 1e-4
        gbn.eval(), obn.eval()
        assert rel_err(gbn(x).F, obn(o).F) < 1e-5
        assert rel_err(ME.MinkowskiELU()(x).F, OME.MinkowskiELU()(o).F) < 1e-6
        xa = x + x
        assert rel_err(xa.F, o.F * 2) < 1e-7
        assert torch.equal(ME.MinkowskiReLU()(x).F.cpu(), torch.relu(o.F))


def test_maxpool_with_normalisation_applied_on_the_fly(dev):
    """b2_maxpool_affine (the stem's InstanceNorm + ReLU folded into the pooling gather) against b2_affine_act followed
    by b2_maxpool on bf16 rows: bit-identical."""
    x, _ = _tensors(dev, 64, batch=3)
    cm, key = x.coordinate_manager, x.coordinate_map_key
    pkey = cm.stride(key, 2)
    km = cm.kernel_map(key, pkey, 2)
    m = cm.get_map(key)
    f = (x.F * 3).to(torch.bfloat16)
    g = torch.Generator().manual_seed(4)
    scale = torch.randn(3, 64, generator=g).to(dev)                 # negative scales included: the max moves rows
    shift = torch.randn(3, 64, generator=g).to(dev)
    for act in (ops.ACT_RELU, ops.ACT_NONE, ops.ACT_ELU):
        want = ops.maxpool(ops.affine_act(f, scale, shift, coords=m.coords, act=act), km)
        got = ops.maxpool_affine(f, km, cm.get_map(pkey).coords, scale, shift, act)
        assert torch.equal(got, want), act


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_instance_norm_statistics_in_one_pass(dev, dtype):
    """b2_instnorm_scale_shift (shifted one-pass sums) against torch: per-scene mean / biased variance folded into
    (scale, shift); an empty scene and a large common offset (the case a naive one-pass sum of squares gets wrong)."""
    g = torch.Generator().manual_seed(8)
    sizes = [5000, 0, 777, 12345]
    off = torch.tensor([0] + list(np.cumsum(sizes)), dtype=torch.int32, device=dev)
    x = (torch.randn(sum(sizes), 64, generator=g) * 0.7 + 40.0).to(dev).to(dtype)
    w = torch.randn(64, generator=g).to(dev)
    b = torch.randn(64, generator=g).to(dev)
    scale, shift = ops.instnorm_scale_shift(x, off, len(sizes), w, b, 1e-8)
    lo = 0
    for i, n in enumerate(sizes):
        if n:
            xs = x[lo:lo + n].double()
            mean, var = xs.mean(0), xs.var(0, unbiased=False)
            want_scale = w.double() / torch.sqrt(var + 1e-8)
            want_shift = b.double() - mean * want_scale
            assert rel_err(scale[i], want_scale.float()) < 1e-4
            assert ((shift[i].double() - want_shift).abs().max() / want_shift.abs().max()).item() < 1e-4
        lo += n
    assert torch.isfinite(scale).all() and torch.isfinite(shift).all()


def test_basic_block_and_union_add(dev):
    from fcaf3d_b200.me.modules.resnet_block import BasicBlock
    x, o = _tensors(dev, 16)
    ods = torch.nn.Sequential(OME.MinkowskiConvolution(16, 32, kernel_size=1, stride=2, dimension=3),
                              OME.MinkowskiBatchNorm(32))
    gds = torch.nn.Sequential(ME.MinkowskiConvolution(16, 32, kernel_size=1, stride=2, dimension=3),
                              ME.MinkowskiBatchNorm(32))
    ob = OME.BasicBlock(16, 32, stride=2, downsample=ods, dimension=3).eval()
    gb = _copy_params(BasicBlock(16, 32, stride=2, downsample=gds, dimension=3), ob).to(dev).eval()
    with torch.no_grad():
        y, ry = gb(x), ob(o)
        assert torch.equal(y.C.cpu(), ry.C) and rel_err(y.F, ry.F) < 1e-4
        # generated children + backbone level: union add (different keys)
        up_o = OME.MinkowskiGenerativeConvolutionTranspose(32, 16, kernel_size=2, stride=2, dimension=3)
        up_g = _copy_params(ME.MinkowskiGenerativeConvolutionTranspose(32, 16, kernel_size=2, stride=2, dimension=3),
                            up_o).to(dev)
        s, rs = x + up_g(y), o + up_o(ry)
        assert torch.equal(s.C.cpu(), rs.C) and rel_err(s.F, rs.F) < 1e-4


@pytest.mark.parametrize("cin,cout,ksize,stride", [(64, 64, 3, 1), (128, 256, 3, 2), (512, 128, 3, 1), (256, 64, 1, 2), (32, 96, 3, 1)])
def test_conv_tf32_tensor_core_path(dev, cin, cout, ksize, stride):
    """b2_conv_tc_tf32 (fp32 rows and weights as TF32 tensor-core operands, fp32 accumulation) with the fused epilogue and a
    residual against the exact fp32 CUDA-core kernel on the same operands.  Stated tolerance: 1e-3 of the output's largest
    magnitude (TF32 keeps 10 mantissa bits of every operand: relative product error <= 2^-10, averaged down by the
    accumulation over K * Cin >= 256 terms), and exactly the sparsity pattern of the exact path (a zero input gives zero)."""
    from util import surface_coords
    c = torch.from_numpy(surface_coords(6000, 2, seed=5)).to(dev)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c)
    cm, key = x0.coordinate_manager, x0.coordinate_map_key
    okey = key if stride == 1 else cm.stride(key, stride)
    km = cm.kernel_map(key, okey, ksize)
    n_in, n_out = cm.size(key), cm.size(okey)
    torch.manual_seed(cin + cout)
    f = torch.randn(n_in, cin, device=dev)
    w = torch.randn(ksize ** 3, cin, cout, device=dev) * 0.05
    wt = w.transpose(1, 2).contiguous()
    scale, shift = torch.rand(cout, device=dev) + .5, torch.randn(cout, device=dev)
    res = torch.randn(n_out, cout, device=dev)
    kw = dict(scale=scale, shift=shift, residual=res, act=ops.ACT_ELU, out_dtype=torch.float32)
    y_tc = ops.sparse_conv(f, w, km, n_out, wt_tf32=wt, **kw)
    y_cc = ops.sparse_conv(f, w, km, n_out, **kw)
    err = ((y_tc - y_cc).abs().max() / y_cc.abs().max()).item()
    assert 0 < err < 1e-3, err                          # > 0: the tensor-core path really ran
    z = ops.sparse_conv(torch.zeros_like(f), w, km, n_out, wt_tf32=wt, out_dtype=torch.float32)
    assert float(z.abs().max()) == 0.0


@pytest.mark.parametrize("cin,cout,stride", [(64, 64, 1), (128, 128, 1), (64, 128, 2), (256, 128, 1)])
def test_persistent_kernel_tf32_is_bit_identical(dev, cin, cout, stride, monkeypatch):
    """The TF32 form of the persistent kernel (fp32 rows, weights and output) against k_conv_tc<..., float> on the same
    layer: equal bit for bit, plain and sparsity-sorted table, fused epilogue with ELU and an fp32 residual."""
    from fcaf3d_b200._lib import lib
    from util import surface_coords
    c = torch.from_numpy(surface_coords(60000, 3, seed=7)).to(dev)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c)
    cm, key = x0.coordinate_manager, x0.coordinate_map_key
    okey = key if stride == 1 else cm.stride(key, stride)
    n_in, n_out = cm.size(key), cm.size(okey)
    torch.manual_seed(3)
    f = torch.randn(n_in, cin, device=dev)
    w = torch.randn(27, cin, cout, device=dev) * 0.05
    wt = w.transpose(1, 2).contiguous()
    scale, shift = torch.rand(cout, device=dev) + .5, torch.randn(cout, device=dev)
    res = torch.randn(n_out, cout, device=dev)
    kw = dict(wt_tf32=wt, scale=scale, shift=shift, residual=res, act=ops.ACT_ELU, out_dtype=torch.float32)
    outs = {}
    for mode in (1, 2):
        old = lib().b2_tc_persist_mode(mode)
        try:
            km = cm.kernel_map(key, okey, 3)
            km.sorted = None
            monkeypatch.setattr(ops, "SORT_MIN_ROWS", 0)
            plain = ops.sparse_conv(f, w, km, n_out, **kw)
            monkeypatch.setattr(ops, "SORT_MIN_ROWS", 1)
            monkeypatch.setattr(ops, "SORT_MIN_USES", 1)
            outs[mode] = (plain, ops.sparse_conv(f, w, km, n_out, **kw))
        finally:
            lib().b2_tc_persist_mode(old)
    assert torch.equal(outs[1][0], outs[2][0]) and torch.equal(outs[1][1], outs[2][1])
    assert torch.equal(outs[2][0], outs[2][1]) and float(outs[2][0].abs().max()) > 0
    y_cc = ops.sparse_conv(f, w, km, n_out, scale=scale, shift=shift, residual=res, act=ops.ACT_ELU, out_dtype=torch.float32)
    assert ((outs[2][0] - y_cc).abs().max() / y_cc.abs().max()).item() < 1e-3
