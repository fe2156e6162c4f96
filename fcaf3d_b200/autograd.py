"""Autograd bindings of the training-path kernels.

Forward / input-gradient run the same sparse-convolution kernels (the input gradient of a convolution is a convolution
over the transposed neighbour table with flipped, transposed weights); the weight gradient is `b2_conv_wgrad`.
Mirrors MinkowskiEngine's MinkowskiConvolutionFunction / MinkowskiLocalPoolingFunction semantics (SURVEY App. A.4,
A.6); everything else of the training graph (BatchNorm1d, ReLU/ELU, indexing) is plain torch autograd, exactly as in
ME, where those modules are torch modules applied to `.F`.
"""
import torch

from . import ops
from ._lib import DT, call, ptr


# With bf16 compute (me.set_compute_dtype(torch.bfloat16)) a tcgen05 convolution hands its output on in bf16, so the
# BatchNorm / activation / residual passes between convolutions move half the bytes and the next convolution needs no
# cast (activation storage as in mixed-precision training; accumulation, BatchNorm statistics, losses and weights stay
# fp32).  False = fp32 activations between the layers (bf16 only inside the convolutions).
BF16_ACTIVATIONS = True


def conv_wgrad(feats, grad_out, kmap, kvol, cin, cout):
    dw = torch.zeros((kvol, cin, cout), dtype=torch.float32, device=feats.device)
    n_out = int(grad_out.shape[0])
    nbr = kmap.nbr if kmap is not None else None
    ld = kmap.ld if kmap is not None else 0
    t0 = ops._prof_start()
    call("b2_conv_wgrad", ptr(feats), DT[feats.dtype], ptr(grad_out), DT[grad_out.dtype], ptr(nbr), ld, n_out, cin, cout,
         kvol, ptr(dw))
    if t0 is not None:
        tc = feats.dtype == torch.bfloat16 and grad_out.dtype == torch.bfloat16 and cin % 64 == 0 and cout % 64 == 0
        pairs = ops.kmap_pairs(kmap, n_out)
        ops._prof_stop(t0, "wgrad_tc" if tc else "wgrad_cc", 2.0 * pairs * cin * cout,
                       feats.element_size() * (feats.shape[0] * cin + n_out * cout) + 4.0 * kvol * (cin * cout + ld))
    return dw


class SparseConvFn(torch.autograd.Function):
    """out[o] = sum_k feats[nbr[k][o]] @ w[k];  w: [K, Cin, Cout] fp32.

    kmap: forward neighbour table (None = identity, K == 1); kmap_t: table of the transposed map (None = identity),
    i.e. for every INPUT row the output rows at coord_in + offset_k' — the pair (i, o) of offset k is found at
    k' = K - 1 - k (odd, centred kernels).
    """

    @staticmethod
    def forward(ctx, feats, w, kmap, kmap_t, n_out, use_bf16, out_dtype=None):
        feats = feats.contiguous()
        wd = w.detach().contiguous()
        kvol, cin, cout = wd.shape
        tc_on = bool(use_bf16) and ops._TC_ENABLED
        tc_fwd = tc_on and cin % 64 == 0 and cout % 32 == 0
        tc_bwd = tc_on and cout % 64 == 0 and cin % 32 == 0 and ctx.needs_input_grad[0]
        x, wt, wt_b = feats, None, None
        if tc_fwd or tc_bwd:
            wt, wt_b = ops.pack_conv_weights(wd.float(), want_fwd=tc_fwd, want_bwd=tc_bwd)   # one launch, both layouts
        if tc_fwd:
            x = ops.cast(feats, torch.bfloat16)
        # bf16 hand-over also for the CUDA-core stem (its [N2, 64] output is the largest activation of the network)
        out_dt = torch.bfloat16 if (BF16_ACTIVATIONS and (tc_fwd or (tc_on and cout % 8 == 0))) else torch.float32
        if out_dtype is not None:             # e.g. the head's logits: fp32 out of the bf16 operands
            out_dt = out_dtype
        out = ops.sparse_conv(x, wd, kmap, n_out, wt_bf16=wt, out_dtype=out_dt)
        ctx.save_for_backward(x, wd, wt_b)
        ctx.kmap, ctx.kmap_t, ctx.n_in = kmap, kmap_t, int(feats.shape[0])
        ctx.in_dtype = feats.dtype
        if out_dt == torch.bfloat16 or out_dtype is not None:
            return out
        return out.to(feats.dtype) if feats.dtype != torch.float32 else out

    @staticmethod
    def backward(ctx, grad_out):
        x, wd, wt_b = ctx.saved_tensors
        kvol, cin, cout = wd.shape
        g = grad_out.contiguous()
        want_x, want_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gb = None
        if (want_x and wt_b is not None) or (want_w and x.dtype == torch.bfloat16):
            gb = ops.cast(g, torch.bfloat16)                              # shared by the input and the weight gradient
        d_feats = d_w = None
        if want_x:
            if wt_b is not None:                                          # [K, Cin, Cout] is already "out x in" for dgrad
                d_feats = ops.sparse_conv(gb, (kvol, cout, cin), ctx.kmap_t, ctx.n_in, wt_bf16=wt_b,
                                          out_dtype=ctx.in_dtype if ctx.in_dtype == torch.bfloat16 else torch.float32)
            else:
                w_back = wd.flip(0).transpose(1, 2).contiguous()          # [K, Cout, Cin]: W[K-1-k']^T
                d_feats = ops.sparse_conv(g.float(), w_back, ctx.kmap_t, ctx.n_in, out_dtype=torch.float32)
            d_feats = d_feats.to(ctx.in_dtype)
        if want_w:
            gw = gb if x.dtype == torch.bfloat16 else (g if g.dtype in DT else g.float())   # fp32 x bf16 is a kernel flavour
            d_w = conv_wgrad(x, gw, ctx.kmap, kvol, cin, cout)
        return d_feats, d_w, None, None, None, None, None


class MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, kmap, n_in):
        out, arg = ops.maxpool(feats.contiguous(), kmap, want_argmax=True)
        ctx.save_for_backward(arg)
        ctx.n_in, ctx.dtype = n_in, feats.dtype
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (arg,) = ctx.saved_tensors
        g = grad_out.contiguous()
        c = int(g.shape[1])
        dx = torch.zeros((ctx.n_in, c), dtype=torch.float32, device=g.device)
        call("b2_maxpool_bwd", ptr(g), DT[g.dtype], ptr(arg), g.shape[0], c, ptr(dx))
        return dx.to(ctx.dtype), None, None


class PruneFn(torch.autograd.Function):
    """row compaction y = x[mask]; backward scatters the gradient rows back (rows that were dropped get zeros)"""

    @staticmethod
    def forward(ctx, feats, m8, pos, n_out):
        ctx.save_for_backward(m8, pos)
        return ops.gather_rows(feats.contiguous(), m8, pos, n_out)

    @staticmethod
    def backward(ctx, grad_out):
        m8, pos = ctx.saved_tensors
        idx = torch.where(m8.bool(), pos, torch.full_like(pos, -1))
        return ops.index_rows(grad_out.contiguous(), idx), None, None, None


# Training-mode BatchNorm (+ residual + ReLU / ELU) on the library's kernels (bn_ops.cu) instead of torch's batch-norm,
# add and activation kernels; False = plain torch modules (what ME itself does).
FUSED_BN = True


class BNActFn(torch.autograd.Function):
    """y = act(batch_norm(x) [+ residual]) over the rows of x [N, C] with batch statistics; running statistics are
    updated in place like nn.BatchNorm1d.  Backward: b2_bn_act_bwd (dx, d weight, d bias, d residual)."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual, bn, act):
        x = x.contiguous()
        n, c = int(x.shape[0]), int(x.shape[1])
        dev = x.device
        stats = torch.empty((4, c), dtype=torch.float32, device=dev)       # mean | invstd | scale | shift
        ws = ops.workspace(ops._lib.lib().b2_bn_ws_bytes(c), dev, slot=3)
        track = bn.track_running_stats and bn.running_mean is not None
        call("b2_bn_train_stats", ptr(x), DT[x.dtype], n, c, ptr(weight.detach().float()) if weight is not None else None,
             ptr(bias.detach().float()) if bias is not None else None, ptr(bn.running_mean) if track else None,
             ptr(bn.running_var) if track else None, ptr(bn.num_batches_tracked) if track else None,
             float(bn.momentum), float(bn.eps), ptr(stats[0]), ptr(stats[1]), ptr(stats[2]), ptr(stats[3]), ptr(ws),
             ws.numel())
        res = residual.contiguous() if residual is not None else None
        y = ops.affine_act(x, stats[2:3], stats[3:4], residual=res, act=act)
        ctx.save_for_backward(x, y if act != ops.ACT_NONE else None, stats, weight)
        ctx.act, ctx.has_res = act, residual is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, stats, weight = ctx.saved_tensors
        n, c = int(x.shape[0]), int(x.shape[1])
        dy = dy.contiguous().to(x.dtype)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if (ctx.has_res and ctx.needs_input_grad[3]) else None
        sums = torch.empty((2, c), dtype=torch.float32, device=x.device)
        ws = ops.workspace(ops._lib.lib().b2_bn_ws_bytes(c), x.device, slot=3)
        call("b2_bn_act_bwd", ptr(x), ptr(dy), ptr(y), DT[x.dtype], n, c, int(ctx.act), ptr(stats[0]), ptr(stats[1]),
             ptr(weight.detach().float()) if weight is not None else None, ptr(sums), ptr(dx), ptr(dres), ptr(ws),
             ws.numel())
        dw = sums[1].to(weight.dtype) if weight is not None and ctx.needs_input_grad[1] else None
        db = sums[0].to(weight.dtype) if weight is not None and ctx.needs_input_grad[2] else None
        return dx, dw, db, dres, None, None


class InstanceNormActFn(torch.autograd.Function):
    """MinkowskiInstanceNorm (+ ReLU) in training: per scene (contiguous row range) and channel statistics, biased
    variance, eps inside the square root, shared affine weight / bias (ME App. A.6; me_resnet.py:22).  Runs the
    BatchNorm kernels of bn_ops.cu once per scene on the row slice: no per-scene torch graph (sum / div / mul / cat and
    their backward accumulations over the largest tensor of the network)."""

    @staticmethod
    def forward(ctx, x, weight, bias, off_host, eps, act):
        x = x.contiguous()
        c = int(x.shape[1])
        nb = len(off_host) - 1
        dev = x.device
        y = torch.empty_like(x)
        stats = torch.zeros((nb, 4, c), dtype=torch.float32, device=dev)   # per scene: mean | invstd | scale | shift
        ws = ops.workspace(ops._lib.lib().b2_bn_ws_bytes(c), dev, slot=3)
        w32, b32 = weight.detach().float().contiguous(), bias.detach().float().contiguous()
        for b in range(nb):
            lo, hi = off_host[b], off_host[b + 1]
            if hi <= lo:
                continue
            xs = x[lo:hi]
            call("b2_bn_train_stats", ptr(xs), DT[x.dtype], hi - lo, c, ptr(w32), ptr(b32), None, None, None, 0.0,
                 float(eps), ptr(stats[b, 0]), ptr(stats[b, 1]), ptr(stats[b, 2]), ptr(stats[b, 3]), ptr(ws), ws.numel())
            ops.affine_act(xs, stats[b, 2:3], stats[b, 3:4], act=act, out=y[lo:hi])
        ctx.save_for_backward(x, y if act != ops.ACT_NONE else None, stats, w32)
        ctx.off_host, ctx.act, ctx.wdtype, ctx.wshape = list(off_host), act, weight.dtype, weight.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, stats, w32 = ctx.saved_tensors
        c = int(x.shape[1])
        nb = len(ctx.off_host) - 1
        dy = dy.contiguous().to(x.dtype)
        dx = torch.zeros_like(x) if any(ctx.off_host[b + 1] <= ctx.off_host[b] for b in range(nb)) else torch.empty_like(x)
        sums = torch.zeros((nb, 2, c), dtype=torch.float32, device=x.device)
        ws = ops.workspace(ops._lib.lib().b2_bn_ws_bytes(c), x.device, slot=3)
        for b in range(nb):
            lo, hi = ctx.off_host[b], ctx.off_host[b + 1]
            if hi <= lo:
                continue
            call("b2_bn_act_bwd", ptr(x[lo:hi]), ptr(dy[lo:hi]), ptr(y[lo:hi]) if y is not None else None, DT[x.dtype],
                 hi - lo, c, int(ctx.act), ptr(stats[b, 0]), ptr(stats[b, 1]), ptr(w32), ptr(sums[b]), ptr(dx[lo:hi]), None,
                 ptr(ws), ws.numel())
        tot = sums.sum(0)
        dw = tot[1].reshape(ctx.wshape).to(ctx.wdtype) if ctx.needs_input_grad[1] else None
        db = tot[0].reshape(ctx.wshape).to(ctx.wdtype) if ctx.needs_input_grad[2] else None
        return dx, dw, db, None, None, None


def bn_act_supported(x, bn):
    return (FUSED_BN and x.is_cuda and x.dim() == 2 and x.shape[0] > 1 and x.shape[1] % 8 == 0 and x.shape[1] <= 2048
            and x.dtype in (torch.float32, torch.bfloat16) and bn.momentum is not None)


class UnionAddFn(torch.autograd.Function):
    """out[dst_a[i]] = a[i]; out[dst_b[j]] += b[j] over the union map (ME A.8: `inputs[i] + x` on different maps);
    backward = two row gathers.  Replaces torch's index_add / gather kernels on bf16 rows (0.2 ms per call)."""

    @staticmethod
    def forward(ctx, a, b, dst_a, dst_b, n):
        a = a.contiguous()
        b = b.contiguous().to(a.dtype)
        out = torch.zeros((n, a.shape[1]), dtype=a.dtype, device=a.device)
        ops.scatter_rows(a, dst_a, out, accumulate=False)
        ops.scatter_rows(b, dst_b, out, accumulate=True)
        ctx.save_for_backward(dst_a, dst_b)
        return out

    @staticmethod
    def backward(ctx, g):
        dst_a, dst_b = ctx.saved_tensors
        g = g.contiguous()
        da = ops.index_rows(g, dst_a) if ctx.needs_input_grad[0] else None
        db = ops.index_rows(g, dst_b) if ctx.needs_input_grad[1] else None
        return da, db, None, None, None


def wants_grad(*tensors):
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)
