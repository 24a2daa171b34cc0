"""Training path of stage 2 (HAT) on libsrb kernels — SURVEY.md §8 row a12, second half of the model.

``hat_forward(stage2, x)`` is differentiable w.r.t. every HAT parameter and w.r.t. its input (the stage-1 output, so
that the gradient reaches RRDBNet).  Forward AND backward run on libsrb operators: tcgen05 convs / linears for the
forward and every data gradient, the split-K weight-gradient kernel, LayerNorm / GELU / window-attention /
overlapping-attention / channel-attention forward+backward kernels; torch allocates, slices, pads and permutes.

Upstream semantics followed (restated in oracle/hat.py from XPixelGroup/HAT hat_arch.py, reached from the reference's
src/architecture.py:51-61,78): HAB = norm1 -> {(S)W-MSA, CAB * 0.01} + shortcut, then norm2 -> MLP + shortcut;
OCAB; RHAG = 6 HAB + OCAB + conv3x3 + residual; conv_first / patch-embed norm / norm / conv_after_body + long skip /
conv_before_upsample + LeakyReLU(0.01) / conv + PixelShuffle(2) / conv_last.  DropPath (stochastic depth, rates
linspace(0, 0.1, 24) on the HAB branches) is a per-sample scale of the branch, drawn from torch's RNG like upstream.

Layout: NHWC bf16 activations and activation gradients ([B,H,W,C], tokens = pixels), fp32 parameter gradients.
The un-fused operators are used on purpose: the fused inference kernels (hab_tc, mlp_tc) keep qkv / hidden activations
on chip, the backward needs them in HBM.
"""
import threading

import torch

from . import _lib as L
from . import ops

_CONV_SCALE = 0.01
_LRELU = 0.01  # nn.LeakyReLU() default slope in conv_before_upsample


def _pad_w(w, cout=None, cin=None):
    co, ci = w.shape[:2]
    out = torch.zeros(cout or co, cin or ci, 3, 3, dtype=torch.float32, device=w.device)
    out[:co, :ci] = w
    return out


def drop_path_masks(B, rate, training, device):
    """upstream drop_path(): keep + rand floor, divided by keep — one factor per sample (None when inactive)."""
    if not training or rate == 0.0:
        return None
    keep = 1.0 - rate
    return (keep + torch.rand(B, device=device)).floor_() / keep  # stays on the device: no sync in the step


def _scale_samples(t, m, plus=None):
    """m[b] * t[b] (+ plus[b]): one pass; m = per-sample factors (device tensor, or a list in tests)"""
    a = torch.as_tensor(m, dtype=torch.float32).to(t.device)
    return ops.axpby_samples(torch.empty_like(t), a, t, 0.0 if plus is None else 1.0, plus)


class _Tape:
    """forward helpers that remember nothing themselves: the caller stores what its backward needs"""

    def __init__(self, P, params):
        self.P = P
        self.params = params  # name -> the nn.Parameter object (carries the packed-weight cache)

    def packed(self, name):
        return ops.packed_weight(self.params[name + '.weight'], self.P[name + '.weight'])

    def conv(self, x, cin, name, out=None, **kw):
        w, b = self.P[name + '.weight'], self.P[name + '.bias']
        if out is None:
            B, H, W = x.shape[:3]
            out = torch.empty(B, H, W, w.shape[0], dtype=torch.bfloat16, device=x.device)
        ops.conv_nhwc(x, cin, self.packed(name), b, out=out, **kw)
        return out


    def conv_gelu(self, x, cin, name):
        """(GELU(conv(x)), conv(x)): the activated output and the bf16 pre-activation the backward needs, one kernel"""
        w, b = self.P[name + '.weight'], self.P[name + '.bias']
        B, H, W = x.shape[:3]
        pre = torch.empty(B, H, W, w.shape[0], dtype=torch.bfloat16, device=x.device)
        act = ops.conv_nhwc_train(x, cin, self.packed(name), b, act=L.ACT_GELU, pre_out=pre)
        return act, pre


    def conv_droppath(self, x, cin, name, m, res, res2=None, res2_scale=None, res2_cs=None):
        """res + m[image] * conv(x) (+ res2_scale * res2): the DropPath-scaled residual branch, and HAB's conv_scale * CAB
        term, in the conv epilogue (no extra pass); m = None: factor 1"""
        w, b = self.P[name + '.weight'], self.P[name + '.bias']
        a = None if m is None else torch.as_tensor(m, dtype=torch.float32).to(x.device)
        cs = res2_cs  # fp32 [B, Cout] per-image channel factors of res2
        if res2 is not None and cs is None:
            cs = torch.full((x.shape[0], w.shape[0]), float(res2_scale), dtype=torch.float32, device=x.device)
        return ops.conv_nhwc_train(x, cin, self.packed(name), b, res1=res, alpha1_samples=a, res2=res2, res2_cs=cs)


class _Grads:
    def __init__(self, P, params):
        self.P = P
        self.params = params
        self.g = {}

    def wgrad(self, name, x, cin, dy, cout, taps):
        dw, db = ops.conv_wgrad(x, cin, dy, cout, taps)
        self.g[name + '.weight'] = dw
        self.g[name + '.bias'] = db

    def dgrad(self, name, dy, out=None, acc=False, gelu_grad_of=None):
        """out (+)= conv^T(dy) through the weights of `name` (3x3 conv or Linear); with ``gelu_grad_of`` (the saved
        pre-activation that fed a GELU producing this layer's input) the result is multiplied by GELU' of it"""
        w = self.P[name + '.weight']
        wd = ops.packed_weight(self.params[name + '.weight'], w, dgrad=True)
        cout, cin = w.shape[0], w.shape[1]
        if out is None:
            B, H, W = dy.shape[:3]
            out = torch.empty(B, H, W, cin, dtype=torch.bfloat16, device=dy.device)
        if gelu_grad_of is not None:
            assert not acc
            return ops.conv_nhwc_train(dy, cout, wd, None, out=out, gelu_grad_of=gelu_grad_of)
        ops.conv_nhwc(dy, cout, wd, None, out=out, res1=out if acc else None)
        return out

    def ln_bwd(self, name, x, g, res=None):
        B, H, W, C = x.shape
        dx, dg, db = ops.layernorm96_bwd(x.view(-1, 96), g.view(-1, 96), self.P[name + '.weight'],
                                         None if res is None else res.view(-1, 96))
        self.g[name + '.weight'] = dg
        self.g[name + '.bias'] = db
        return dx.view(B, H, W, C)


def _ln(P, name, x):
    B, H, W, C = x.shape
    return ops.layernorm96(x.view(-1, 96), P[name + '.weight'], P[name + '.bias']).view(B, H, W, C)


# ---------------------------------------------------------------------------------------------------------------------
# MLP tail shared by HAB and OCAB:  x2 = x1 + m * fc2(gelu(fc1(norm2(x1))))
# ---------------------------------------------------------------------------------------------------------------------
def _mlp_fwd(T, pre, x1, m):
    P = T.P
    x1n = _ln(P, pre + 'norm2', x1)
    ha, hp = T.conv_gelu(x1n, 96, pre + 'mlp.fc1')
    if m is None:
        x2 = T.conv(ha, 192, pre + 'mlp.fc2', res1=x1)
    else:
        x2 = T.conv_droppath(ha, 192, pre + 'mlp.fc2', m, x1)
    return x2, (x1n, hp, ha)  # ha is kept: recomputing it costs a full HBM pass over [tokens, 192]


def _mlp_bwd(G, pre, x1, saved, m, g2):
    """g2 = dL/dx2 -> dL/dx1 (identity path included)"""
    x1n, hp, ha = saved
    gb = g2 if m is None else _scale_samples(g2, m)
    G.wgrad(pre + 'mlp.fc2', ha, 192, gb, 96, 1)
    dhp = G.dgrad(pre + 'mlp.fc2', gb, gelu_grad_of=hp)
    G.wgrad(pre + 'mlp.fc1', x1n, 96, dhp, 192, 1)
    dx1n = G.dgrad(pre + 'mlp.fc1', dhp)
    return G.ln_bwd(pre + 'norm2', x1, dx1n, res=g2)


# ---------------------------------------------------------------------------------------------------------------------
# HAB
# ---------------------------------------------------------------------------------------------------------------------
def _hab_fwd(T, pre, x, shift, rpi, m1, m2):
    P = T.P
    B, H, W, _ = x.shape
    xn = _ln(P, pre + 'norm1', x)
    qkv = T.conv(xn, 96, pre + 'attn.qkv')
    bias = ops.gather_rel_bias(P[pre + 'attn.relative_position_bias_table'], rpi)
    ao = ops.window_attention(qkv, bias, shift)
    c1, c1p = T.conv_gelu(xn, 96, pre + 'conv_block.cab.0')
    c2 = T.conv(c1, 32, pre + 'conv_block.cab.2')
    ca_pre = pre + 'conv_block.cab.3.attention.'
    _, ca_saved = ops.channel_attention(c2.view(B, H * W, 96), P[ca_pre + '1.weight'].view(3, 96), P[ca_pre + '1.bias'],
                                        P[ca_pre + '3.weight'].view(96, 3), P[ca_pre + '3.bias'], want_out=False)
    # x1 = shortcut + DropPath(proj(attention)) + conv_scale * CAB: all three terms in the proj epilogue; the channel
    # attention scale s[b][c] rides in the per-image channel factor of res2 (no scaled copy of c2 is ever written)
    x1 = T.conv_droppath(ao, 96, pre + 'attn.proj', m1, x, res2=c2, res2_cs=ca_saved[2] * _CONV_SCALE)
    x2, mlp_saved = _mlp_fwd(T, pre, x1, m2)
    return x2, (x, xn, qkv, bias, ao, c1p, c1, c2, ca_saved, x1, mlp_saved)


def _hab_bwd(G, pre, saved, shift, rpi, m1, m2, g2):
    P = G.P
    x, xn, qkv, bias, ao, c1p, c1, c2, ca_saved, x1, mlp_saved = saved
    B, H, W, _ = x.shape
    g1 = _mlp_bwd(G, pre, x1, mlp_saved, m2, g2)                       # dL/dx1
    # ---- attention branch
    gb = g1 if m1 is None else _scale_samples(g1, m1)
    G.wgrad(pre + 'attn.proj', ao, 96, gb, 96, 1)
    dao = G.dgrad(pre + 'attn.proj', gb)
    dqkv, dtab = ops.window_attention_bwd(qkv, bias, rpi, dao, shift)
    G.g[pre + 'attn.relative_position_bias_table'] = dtab
    G.wgrad(pre + 'attn.qkv', xn, 96, dqkv, 288, 1)
    dxn = G.dgrad(pre + 'attn.qkv', dqkv)
    # ---- CAB branch (scaled by conv_scale)
    ca_pre = pre + 'conv_block.cab.3.attention.'
    dc2, dw1, db1, dw2, db2 = ops.channel_attention_bwd(c2.view(B, H * W, 96), g1.view(B, H * W, 96),
                                                        P[ca_pre + '1.weight'].view(3, 96),
                                                        P[ca_pre + '3.weight'].view(96, 3), ca_saved, g_scale=_CONV_SCALE)
    G.g[ca_pre + '1.weight'] = dw1.view(3, 96, 1, 1)
    G.g[ca_pre + '1.bias'] = db1
    G.g[ca_pre + '3.weight'] = dw2.view(96, 3, 1, 1)
    G.g[ca_pre + '3.bias'] = db2
    dc2 = dc2.view(B, H, W, 96)
    G.wgrad(pre + 'conv_block.cab.2', c1, 32, dc2, 96, 9)
    dc1p = G.dgrad(pre + 'conv_block.cab.2', dc2, gelu_grad_of=c1p)
    G.wgrad(pre + 'conv_block.cab.0', xn, 96, dc1p, 32, 9)
    G.dgrad(pre + 'conv_block.cab.0', dc1p, out=dxn, acc=True)         # both consumers of norm1's output
    return G.ln_bwd(pre + 'norm1', x, dxn, res=g1)


# ---------------------------------------------------------------------------------------------------------------------
# OCAB
# ---------------------------------------------------------------------------------------------------------------------
def _ocab_fwd(T, pre, x, rpi):
    P = T.P
    xn = _ln(P, pre + 'norm1', x)
    qkv = T.conv(xn, 96, pre + 'qkv')
    bias = ops.gather_rel_bias_oca(P[pre + 'relative_position_bias_table'], rpi)
    ao = ops.overlap_attention(qkv, bias)
    x1 = T.conv(ao, 96, pre + 'proj', res1=x)
    x2, mlp_saved = _mlp_fwd(T, pre, x1, None)
    return x2, (x, xn, qkv, bias, ao, x1, mlp_saved)


def _ocab_bwd(G, pre, saved, rpi, g2):
    x, xn, qkv, bias, ao, x1, mlp_saved = saved
    g1 = _mlp_bwd(G, pre, x1, mlp_saved, None, g2)
    G.wgrad(pre + 'proj', ao, 96, g1, 96, 1)
    dao = G.dgrad(pre + 'proj', g1)
    dqkv, dtab = ops.overlap_attention_bwd(qkv, bias, rpi, dao)
    G.g[pre + 'relative_position_bias_table'] = dtab
    G.wgrad(pre + 'qkv', xn, 96, dqkv, 288, 1)
    dxn = G.dgrad(pre + 'qkv', dqkv)
    return G.ln_bwd(pre + 'norm1', x, dxn, res=g1)


# ---------------------------------------------------------------------------------------------------------------------
# whole network
# ---------------------------------------------------------------------------------------------------------------------
class _HATFn(torch.autograd.Function):
    _tls = threading.local()  # .cfg = (names, rpi_sa, rpi_oca, masks, debug) for the call that follows

    @staticmethod
    def forward(ctx, x, *params):
        if x.device.type != 'cuda':
            raise RuntimeError('libsrb training path runs on CUDA only; no CPU fallback')
        names, rpi_sa, rpi_oca, masks, debug = _HATFn._tls.cfg
        _HATFn._tls.cfg = None
        P = {n: p.detach().float().contiguous() for n, p in zip(names, params)}
        B, _, H, W = x.shape
        if H % 16 or W % 32:
            raise ValueError('hat_forward: H must be a multiple of 16 and W of 32 (tile and weight-gradient chunking)')
        dev = x.device
        PO = dict(zip(names, params))
        T = _Tape(P, PO)
        x16 = torch.zeros(B, H, W, 16, dtype=torch.bfloat16, device=dev)
        x16[..., 0] = x.detach()[:, 0]
        wf = _pad_w(P['conv_first.weight'], cin=16)
        f = torch.empty(B, H, W, 96, dtype=torch.bfloat16, device=dev)
        ops.conv_nhwc(x16, 16, ops.pack_conv_weight(wf), P['conv_first.bias'], out=f)
        t = _ln(P, 'patch_embed.norm', f)
        saved_layers = []
        for l in range(4):
            t_in = t
            blocks = []
            for j in range(6):
                pre = f'layers.{l}.residual_group.blocks.{j}.'
                m1, m2 = masks[l * 6 + j]
                t, sv = _hab_fwd(T, pre, t, 4 if j % 2 else 0, rpi_sa, m1, m2)
                blocks.append(sv)
            t, osv = _ocab_fwd(T, f'layers.{l}.residual_group.overlap_attn.', t, rpi_oca)
            t_oc = t
            t = T.conv(t_oc, 96, f'layers.{l}.conv', res1=t_in)
            saved_layers.append((blocks, osv, t_oc))
        tN = t
        tn = _ln(P, 'norm', tN)
        r = T.conv(tn, 96, 'conv_after_body', res1=f)
        u = T.conv(r, 96, 'conv_before_upsample.0', act=L.ACT_LRELU, slope=_LRELU)
        s = torch.empty(B, 2 * H, 2 * W, 64, dtype=torch.bfloat16, device=dev)
        ops.conv_nhwc(u, 64, T.packed('upsample.0'), P['upsample.0.bias'], out=s, pixel_shuffle=True)
        y = ops.conv3x3_cout1(s, P['conv_last.weight'], P['conv_last.bias'])  # fp32 [B,2H,2W]
        if debug is not None:  # test hook: the stored LeakyReLU output (its sign pattern decides the backward)
            debug['u'] = u
        ctx.params = PO
        ctx.saved = (P, names, rpi_sa, rpi_oca, masks, x16, f, saved_layers, tN, tn, r, u, s)
        ctx.geom = (B, H, W)
        ctx.need_dx = ctx.needs_input_grad[0]
        return y.unsqueeze(1)

    @staticmethod
    def backward(ctx, gy):
        P, names, rpi_sa, rpi_oca, masks, x16, f, saved_layers, tN, tn, r, u, s = ctx.saved
        B, H, W = ctx.geom
        dev = gy.device
        bf = dict(dtype=torch.bfloat16, device=dev)
        G = _Grads(P, ctx.params)
        # ---- conv_last (64 -> 1): gradient padded to 16 channels so that the tensor-core kernels apply
        g16 = torch.zeros(B, 2 * H, 2 * W, 16, **bf)
        g16[..., 0] = gy.detach()[:, 0]
        dw, db = ops.conv_wgrad(s, 64, g16, 8, 9)
        G.g['conv_last.weight'] = dw[:1].contiguous()
        G.g['conv_last.bias'] = db[:1].contiguous()
        gs = torch.empty(B, 2 * H, 2 * W, 64, **bf)
        ops.conv_nhwc(g16, 16, ops.pack_conv_weight_dgrad(_pad_w(P['conv_last.weight'], cout=16)), None, out=gs)
        # ---- PixelShuffle(2) backward = pixel-unshuffle (pure data movement): channel c*4 + dy*2 + dx
        gps = gs.view(B, H, 2, W, 2, 64).permute(0, 1, 3, 5, 2, 4).reshape(B, H, W, 256).contiguous()
        G.wgrad('upsample.0', u, 64, gps, 256, 9)
        gu = G.dgrad('upsample.0', gps)
        ops.lrelu_bwd_(gu, u, _LRELU)
        G.wgrad('conv_before_upsample.0', r, 96, gu, 64, 9)
        gr = G.dgrad('conv_before_upsample.0', gu)           # dL/dr;  r = conv_after_body(norm(tN)) + f
        G.wgrad('conv_after_body', tn, 96, gr, 96, 9)
        gtn = G.dgrad('conv_after_body', gr)
        g = G.ln_bwd('norm', tN, gtn)
        for l in reversed(range(4)):
            blocks, osv, t_oc = saved_layers[l]
            gl = g                                            # dL/d(layer output) = conv(t_oc) + t_in
            G.wgrad(f'layers.{l}.conv', t_oc, 96, gl, 96, 9)
            g = G.dgrad(f'layers.{l}.conv', gl)
            g = _ocab_bwd(G, f'layers.{l}.residual_group.overlap_attn.', osv, rpi_oca, g)
            for j in reversed(range(6)):
                m1, m2 = masks[l * 6 + j]
                g = _hab_bwd(G, f'layers.{l}.residual_group.blocks.{j}.', blocks[j], 4 if j % 2 else 0, rpi_sa, m1, m2, g)
                blocks[j] = None                              # release the block's saved activations
            g = ops.axpby(g, 1.0, g, 1.0, gl)                 # + the RHAG residual
        gf = G.ln_bwd('patch_embed.norm', f, g, res=gr)       # f feeds the patch-embed norm and the long skip
        dw, db = ops.conv_wgrad(x16, 16, gf, 96, 9)
        G.g['conv_first.weight'] = dw[:, :1].contiguous()
        G.g['conv_first.bias'] = db
        gx = None
        if ctx.need_dx:
            g32 = torch.empty(B, H, W, 32, **bf)
            ops.conv_nhwc(gf, 96, ops.pack_conv_weight_dgrad(_pad_w(P['conv_first.weight'], cin=32)), None, out=g32)
            gx = g32[..., 0].float().unsqueeze(1)
        ctx.saved = None
        return (gx,) + tuple(G.g[n] for n in names)


def hat_forward(stage2, x, drop_path_rate=0.1, training=None, masks=None, debug=None):
    """Differentiable HAT forward on libsrb: x [B,1,H,W] float -> [B,1,2H,2W] fp32.
    ``stage2``: the ``stage2`` sub-module of ``HybridSuperResolutionModel`` (parameter holder, XPixelGroup/HAT key layout).
    ``drop_path_rate``: upstream's un-passed constructor default 0.1 (stochastic depth, active when ``training``;
    ``training`` defaults to ``stage2.training``)."""
    if training is None:
        training = stage2.training
    from .architecture import named_params
    named = named_params(stage2)  # also valid on nn.DataParallel replicas
    names = [n for n, _ in named]
    B = x.shape[0]
    rates = torch.linspace(0, drop_path_rate, 24).tolist()
    if masks is None:  # (attention-branch factors, MLP-branch factors) per HAB, each a list of B floats or None
        masks = [(drop_path_masks(B, rt, training, x.device), drop_path_masks(B, rt, training, x.device)) for rt in rates]
    _HATFn._tls.cfg = (names, stage2.relative_position_index_SA, stage2.relative_position_index_OCA, masks, debug)
    return _HATFn.apply(x, *[p for _, p in named])
