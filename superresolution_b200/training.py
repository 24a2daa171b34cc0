"""Training path of stage 1 (RRDBNet) on libsrb kernels — SURVEY.md §8 row a12, first half of the model.

The reference trains ``HybridSuperResolutionModel`` through PyTorch autograd (``scripts/Modello_supporto.py:125-141``);
with HAT missing it is exactly this network + an interpolation (``src/architecture.py:35-47,80-81``, has_stage2=False).
``rrdbnet_forward(stage1, x)`` is differentiable w.r.t. every stage-1 parameter: forward AND backward run on the libsrb
operators (tcgen05 convs for the forward and the data gradients, the split-K weight-gradient kernel, LeakyReLU /
upsample / residual backward kernels); torch only allocates, slices, pads and permutes.

Layout: NHWC bf16.  One ``[B,h,w,192]`` buffer per residual dense block holds ``[x|x1|x2|x3|x4]`` (kept for the
backward: it is the input of every conv of the block and the sign source of every LeakyReLU); the residual trunk and
its gradient are fp32, the per-block gradient buffer ``G[B,h,w,192]`` is bf16 (dense-block concat backward = in-place
accumulation into channel slices through the conv epilogue's ``+ res1``).

``hybrid_forward(model, x)`` chains it with the blur layers, the stage-2 path (training_hat.py) and the bicubic resize
branch into the differentiable forward of ``HybridSuperResolutionModel`` (src/architecture.py:74-82); ``train_step`` is
the optimizer step of ``scripts/Modello_supporto.py:125-141`` over it.
"""
import threading

import torch
import torch.nn.functional as F

from . import _lib as L
from . import ops

_LRELU = 0.2


def _names():
    yield 'conv_first'
    for i in range(23):
        for r in (1, 2, 3):
            for k in (1, 2, 3, 4, 5):
                yield f'body.{i}.rdb{r}.conv{k}'
    for n in ('conv_body', 'conv_up1', 'conv_up2', 'conv_hr', 'conv_last'):
        yield n


CONV_NAMES = list(_names())  # parameter order of RRDBNet = state-dict order (weight, bias per conv)


def _pad_w(w, cout=None, cin=None):
    """zero-pad a conv weight [Cout,Cin,3,3] to the channel counts the tensor-core kernels accept"""
    co, ci = w.shape[:2]
    out = torch.zeros(cout or co, cin or ci, 3, 3, dtype=torch.float32, device=w.device)
    out[:co, :ci] = w
    return out


class _RRDBNetFn(torch.autograd.Function):
    _tls = threading.local()  # .direct: set by rrdbnet_forward(accumulate_into_grad=True) for the call that follows
                              # (thread-local: nn.DataParallel drives one replica per thread)

    @staticmethod
    def forward(ctx, x, *params):
        if x.device.type != 'cuda':
            raise RuntimeError('libsrb training path runs on CUDA only; no CPU fallback')
        P = {n: (params[2 * i].detach().float(), params[2 * i + 1].detach().float()) for i, n in enumerate(CONV_NAMES)}
        B, _, H, W = x.shape
        h, w = H // 2, W // 2
        if h % 16 or w % 32:
            raise ValueError('rrdbnet_forward: H must be a multiple of 32 and W of 64 (tile and weight-gradient chunking)')
        dev = x.device
        bf = dict(dtype=torch.bfloat16, device=dev)
        # pixel_unshuffle(2) -> NHWC, padded to 16 channels (data movement only)
        u16 = torch.zeros(B, h, w, 16, **bf)
        u16[..., :4] = F.pixel_unshuffle(x.detach().float(), 2).permute(0, 2, 3, 1)
        D = [torch.empty(B, h, w, 192, **bf) for _ in range(69)]
        t_rrdb = torch.empty(B, h, w, 64, dtype=torch.float32, device=dev)  # fp32 trunk: RRDB input
        t_a = torch.empty_like(t_rrdb)
        t_b = torch.empty_like(t_rrdb)
        wf, bfirst = P['conv_first']
        ops.conv_nhwc(u16, 16, ops.pack_conv_weight(_pad_w(wf, cin=16)), bfirst, out=D[0][..., :64], out32=t_rrdb)
        f0_32 = t_rrdb.clone()
        body_out = torch.empty(B, h, w, 64, **bf)
        for i in range(23):
            t_in = t_rrdb  # trunk entering the current RDB
            for r in range(3):
                Dn = D[3 * i + r]
                for k in range(1, 5):
                    wk, bk = P[f'body.{i}.rdb{r + 1}.conv{k}']
                    ops.conv_nhwc(Dn, 32 + 32 * k, ops.pack_conv_weight(wk), bk, out=Dn[..., 32 + 32 * k:64 + 32 * k],
                                  act=L.ACT_LRELU, slope=_LRELU)
                w5, b5 = P[f'body.{i}.rdb{r + 1}.conv5']
                last = r == 2
                nxt = body_out if (i == 22 and last) else D[3 * i + r + 1][..., :64]
                t_out = t_b if t_in is t_a else t_a
                if last:  # (x5*0.2 + rdb_in)*0.2 + rrdb_in
                    ops.conv_nhwc(Dn, 192, ops.pack_conv_weight(w5), b5, out=nxt, out32=t_out, alpha1=0.2, res1=t_in,
                                  alpha2=0.2, res2=t_rrdb)
                    t_rrdb.copy_(t_out)
                else:     # x5*0.2 + rdb_in
                    ops.conv_nhwc(Dn, 192, ops.pack_conv_weight(w5), b5, out=nxt, out32=t_out, alpha1=0.2, res1=t_in)
                    t_in = t_out
        wb, bb = P['conv_body']
        feat = torch.empty(B, h, w, 64, **bf)
        ops.conv_nhwc(body_out, 64, ops.pack_conv_weight(wb), bb, out=feat, res1=f0_32)  # feat = f0 + conv_body(body)
        U1 = ops.upsample2x_nearest(feat)
        A1 = torch.empty(B, 2 * h, 2 * w, 64, **bf)
        ops.conv_nhwc(U1, 64, ops.pack_conv_weight(P['conv_up1'][0]), P['conv_up1'][1], out=A1, act=L.ACT_LRELU, slope=_LRELU)
        U2 = ops.upsample2x_nearest(A1)
        A2 = torch.empty(B, 4 * h, 4 * w, 64, **bf)
        ops.conv_nhwc(U2, 64, ops.pack_conv_weight(P['conv_up2'][0]), P['conv_up2'][1], out=A2, act=L.ACT_LRELU, slope=_LRELU)
        A3 = torch.empty_like(A2)
        ops.conv_nhwc(A2, 64, ops.pack_conv_weight(P['conv_hr'][0]), P['conv_hr'][1], out=A3, act=L.ACT_LRELU, slope=_LRELU)
        y = ops.conv3x3_cout1(A3, P['conv_last'][0], P['conv_last'][1])  # fp32 [B,4h,4w]
        ctx.saved = (P, u16, D, body_out, U1, A1, U2, A2, A3)
        ctx.geom = (B, h, w)
        ctx.direct = getattr(_RRDBNetFn._tls, 'direct', None) or {}
        _RRDBNetFn._tls.direct = None
        return y.unsqueeze(1)

    @staticmethod
    def backward(ctx, gy):
        P, u16, D, body_out, U1, A1, U2, A2, A3 = ctx.saved
        B, h, w = ctx.geom
        dev = gy.device
        bf = dict(dtype=torch.bfloat16, device=dev)
        grads = {}

        def wgrad(name, x, cin, dy, cout, keep_out=None, keep_in=None):
            tgt = ctx.direct.get(name) if not (keep_out or keep_in) else None
            if tgt is not None:  # accumulate straight into the parameters' .grad buffers (no autograd add kernels)
                ops.conv_wgrad(x, cin, dy, cout, 9, dw=tgt[0], db=tgt[1], accumulate=True)
                grads[name] = (None, None)
                return
            dw, db = ops.conv_wgrad(x, cin, dy, cout, 9)
            grads[name] = (dw[:keep_out, :keep_in].contiguous() if (keep_out or keep_in) else dw,
                           db[:keep_out].contiguous() if keep_out else db)

        def dgrad(name, dy, cout, out, cin_rows=None, acc=False, pad_out=None):
            """out (+)= conv^T(dy) for the input channels [0, cin_rows) of conv `name`"""
            wq = P[name][0]
            if cin_rows is not None:
                wq = wq[:, :cin_rows]
            if pad_out:
                wq = _pad_w(wq, cout=pad_out)
            ops.conv_nhwc(dy, pad_out or cout, ops.pack_conv_weight_dgrad(wq.contiguous()), None, out=out,
                          res1=out if acc else None)

        # ---- conv_last (64 -> 1): dy padded to 16 channels so that the tensor-core kernels apply
        g16 = torch.zeros(B, 4 * h, 4 * w, 16, **bf)
        g16[..., 0] = gy.detach()[:, 0]
        wgrad('conv_last', A3, 64, g16, 8, keep_out=1)
        g = torch.empty(B, 4 * h, 4 * w, 64, **bf)
        dgrad('conv_last', g16, 1, g, pad_out=16)
        ops.lrelu_bwd_(g, A3, _LRELU)
        # ---- conv_hr
        wgrad('conv_hr', A2, 64, g, 64)
        g2 = torch.empty_like(g)
        dgrad('conv_hr', g, 64, g2)
        ops.lrelu_bwd_(g2, A2, _LRELU)
        # ---- conv_up2 on U2 = nearest2x(A1)
        wgrad('conv_up2', U2, 64, g2, 64)
        dgrad('conv_up2', g2, 64, g)
        gA1 = ops.upsample2x_bwd(g)
        ops.lrelu_bwd_(gA1, A1, _LRELU)
        # ---- conv_up1 on U1 = nearest2x(feat)
        wgrad('conv_up1', U1, 64, gA1, 64)
        gU1 = torch.empty(B, 2 * h, 2 * w, 64, **bf)
        dgrad('conv_up1', gA1, 64, gU1)
        gfeat = ops.upsample2x_bwd(gU1)                      # bf16 [B,h,w,64]: gradient of feat = f0 + conv_body(body)
        gf0 = torch.empty(B, h, w, 64, dtype=torch.float32, device=dev)
        ops.axpby(gf0, 1.0, gfeat)                           # fp32 accumulator of d/d f0 (long skip)
        # ---- conv_body
        wgrad('conv_body', body_out, 64, gfeat, 64)
        gb = torch.empty(B, h, w, 64, **bf)
        dgrad('conv_body', gfeat, 64, gb)
        gR = torch.empty(B, h, w, 64, dtype=torch.float32, device=dev)  # fp32 trunk gradient: RRDB output
        ops.axpby(gR, 1.0, gb)
        G = torch.empty(B, h, w, 192, **bf)
        g5 = torch.empty(B, h, w, 64, **bf)
        gO = torch.empty_like(gR)
        for i in reversed(range(23)):
            # out = rdb3_out * 0.2 + rrdb_in
            ops.axpby(gO, 0.2, gR)                           # gradient of rdb3's output
            for r in reversed(range(3)):
                Dn = D[3 * i + r]
                pre = f'body.{i}.rdb{r + 1}.conv'
                ops.axpby(g5, 0.2, gO)                       # x5 * 0.2 + x  ->  d/d x5
                wgrad(pre + '5', Dn, 192, g5, 64)
                dgrad(pre + '5', g5, 64, G)                  # G[0:192) = conv5^T(g5)
                for k in (4, 3, 2, 1):
                    gk = G[..., 32 + 32 * k:64 + 32 * k]
                    ops.lrelu_bwd_(gk, Dn[..., 32 + 32 * k:64 + 32 * k], _LRELU)
                    wgrad(pre + str(k), Dn, 32 + 32 * k, gk, 32)
                    dgrad(pre + str(k), gk, 32, G[..., :32 + 32 * k], acc=True)
                ops.axpby(gO, 1.0, gO, 1.0, G[..., :64])     # identity path + dense-block path -> d/d rdb input
            ops.axpby(gR, 1.0, gR, 1.0, gO)                  # rrdb_in: identity + chain
        # ---- conv_first: feat0 feeds the body (gR) and the long skip (gf0)
        ops.axpby(gf0, 1.0, gf0, 1.0, gR)
        gf0b = torch.empty(B, h, w, 64, **bf)
        ops.axpby(gf0b, 1.0, gf0)
        wgrad('conv_first', u16, 16, gf0b, 64, keep_in=4)
        out = [None]
        for n in CONV_NAMES:
            out += list(grads[n])
        return tuple(out)


def rrdbnet_forward(stage1, x, accumulate_into_grad=False):
    """Differentiable RRDBNet forward on libsrb: x [B,1,H,W] float in [0,1] -> [B,1,2H,2W] fp32.
    ``stage1``: the ``stage1`` sub-module of ``HybridSuperResolutionModel`` (parameter holder, BasicSR key layout).
    ``accumulate_into_grad``: the weight-gradient kernels add straight into existing fp32 ``p.grad`` buffers (e.g. the
    flat-gradient views of ``FusedAdamW``) instead of returning tensors for autograd to add — same result, ~700 fewer
    element-wise launches per step; parameters without a ``.grad`` buffer take the autograd route."""
    from .architecture import named_params
    sd = dict(named_params(stage1))  # also valid on nn.DataParallel replicas
    params = []
    direct = {}
    for n in CONV_NAMES:
        w, b = sd[n + '.weight'], sd[n + '.bias']
        params += [w, b]
        if accumulate_into_grad and w.grad is not None and b.grad is not None and w.grad.is_contiguous() \
                and w.grad.dtype == torch.float32 and b.grad.dtype == torch.float32:
            direct[n] = (w.grad, b.grad)
    _RRDBNetFn._tls.direct = direct
    return _RRDBNetFn.apply(x, *params)


# ---------------------------------------------------------------------------------------------------------------------
# Whole-model training path: reference src/architecture.py:74-82 made differentiable on libsrb operators.
class _BlurFn(torch.autograd.Function):
    """AntiCheckerboardLayer: zero-padded symmetric binomial operator -> its own transpose."""

    @staticmethod
    def forward(ctx, x, mode):
        ctx.mode = mode
        return ops.blur(x.detach().float(), mode)

    @staticmethod
    def backward(ctx, g):
        return ops.blur(g.float(), ctx.mode), None


class _BicubicFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, Ho, Wo):
        ctx.in_hw = tuple(x.shape[-2:])
        return ops.bicubic_resize(x.detach().float(), Ho, Wo)

    @staticmethod
    def backward(ctx, g):
        return ops.bicubic_resize_bwd(g.float(), *ctx.in_hw), None, None


def _head(model, x, accumulate_into_grad):
    """stage1 -> s1 (src/architecture.py:75-76)"""
    y = rrdbnet_forward(model.stage1, x, accumulate_into_grad=accumulate_into_grad)
    mode = getattr(model.s1, 'mode', None)  # nn.Identity when smoothing == 'none'
    return _BlurFn.apply(y, mode) if mode is not None else y


def _tail(model, y, drop_path_rate, masks=None, debug=None):
    """stage2 -> s2 -> bicubic resize when the shape differs -> sf (src/architecture.py:77-82)"""
    from .training_hat import hat_forward
    mode = getattr(model.s1, 'mode', None)
    y = hat_forward(model.stage2, y, drop_path_rate=drop_path_rate, training=model.training, masks=masks, debug=debug)
    if mode is not None:
        y = _BlurFn.apply(y, mode)
    if y.shape[-1] != model.output_size:
        y = _BicubicFn.apply(y, model.output_size, model.output_size)
    if mode is not None:
        y = _BlurFn.apply(y, 'light')
    return y


def hybrid_forward(model, x, accumulate_into_grad=False, drop_path_rate=0.1, masks=None, debug=None):
    """Differentiable ``HybridSuperResolutionModel.forward``: stage1 -> s1 -> stage2 -> s2 -> (bicubic resize to
    output_size when the shape differs) -> sf, exactly the order of src/architecture.py:74-82.  x: [B,1,H,W] float in
    [0,1] with H % 32 == 0 and W % 64 == 0.  Returns fp32 [B,1,output_size,output_size] (or [B,1,4H,4W]).
    ``masks`` / ``debug``: see training_hat.hat_forward (given DropPath factors / captured intermediates; tests)."""
    return _tail(model, _head(model, x, accumulate_into_grad), drop_path_rate, masks, debug)


def train_step(model, optimizer, criterion, lr, hr, micro_batch=16, max_norm=1.0, world=1, group=None, grad_scale=1.0,
               drop_path_rate=0.1):
    """One optimizer step of scripts/Modello_supporto.py:125-141 on libsrb: forward + loss + backward over the batch,
    gradient SUM all-reduce over ``world`` ranks (mean folded into the update), global-norm clip at ``max_norm`` and
    AdamW — the last three fused in ``FusedAdamW.step``.

    Memory plan: stage 1 (RRDBNet, ~0.11 GB of saved activations per patch) runs forward and backward on the WHOLE batch
    — its 700 small launches are latency-bound below ~16 patches; stage 2 (HAT, ~5 GB per patch: it does not fit
    180 GB at 32 patches) runs in ``micro_batch``-sized slices of the stage-1 output, each slice forward + loss +
    backward at once, its input gradient collected into the stage-1 output gradient.  Without batch statistics the sum
    of the slice gradients IS the batch gradient; parameter gradients accumulate in the optimizer's flat fp32 buffer.
    ``model`` must be in train() mode, on CUDA; ``optimizer`` a ``FusedAdamW`` over ``model.parameters()``.
    Returns the batch loss (0-d tensor, no sync)."""
    from .optim import allreduce_gradients_start, allreduce_gradients_wait
    B = lr.shape[0]
    mb = micro_batch or B
    optimizer.zero_grad()
    y1 = _head(model, lr, True)
    g1 = torch.empty_like(y1)
    total = None
    with ops.weight_pack_cache():  # the parameters do not change before optimizer.step: pack each weight once
        for s in range(0, B, mb):
            e = min(B, s + mb)
            ys = y1[s:e].detach().requires_grad_(True)
            loss, _ = criterion(_tail(model, ys, drop_path_rate), hr[s:e])
            (loss * (grad_scale * (e - s) / B)).backward()  # mean over the whole batch = weighted sum of slice means
            g1[s:e] = ys.grad
            total = loss.detach() * ((e - s) / B) if total is None else total + loss.detach() * ((e - s) / B)
    works, cut = [], 0
    if world > 1:
        # every stage-2 gradient is final here (the stage-1 backward below does not touch them): their all-reduce
        # starts now and runs under the stage-1 backward.  model.parameters() lists stage 1 first, so stage 2 is the
        # tail of the flat buffer.
        optimizer._gather_stray_grads()  # (no-op on this path: the autograd functions accumulate into the flat views)
        cut = stage2_offset(model, optimizer)
        works = allreduce_gradients_start(optimizer.flat_grad, group, lo=cut)
    y1.backward(g1)
    if world > 1:
        optimizer._gather_stray_grads()
        works += allreduce_gradients_start(optimizer.flat_grad, group, hi=cut)
        allreduce_gradients_wait(works)
    optimizer.step(grad_scale=grad_scale, max_norm=max_norm, world=world)
    return total


def stage2_offset(model, optimizer):
    """flat-buffer offset of the first stage-2 parameter when stage 2 forms the tail of the optimizer's parameter list
    (the order of ``HybridSuperResolutionModel.parameters()``); otherwise the buffer's length = nothing is reduced early"""
    m = getattr(model, 'module', model)
    s2 = {id(p) for p in m.stage2.parameters()}
    ps = optimizer.param_groups[0]['params']
    first = next((i for i, p in enumerate(ps) if id(p) in s2), None)
    if first is None or any(id(p) not in s2 for p in ps[first:]):
        return optimizer.numel
    return optimizer.offset_of(ps[first])
