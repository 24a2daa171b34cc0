"""Granular operator wrappers over the C ABI (used by the parity tests; the model forward uses srb_forward)."""
import math

import torch

from . import _lib as L


_PACK_SCOPE = 0  # id of the active weight_pack_cache() scope, 0 = none: every call packs afresh


class weight_pack_cache:
    """``with ops.weight_pack_cache():`` — inside the scope ``packed_weight`` keeps each parameter's packed bf16 image
    on the parameter object, so that the micro-batches of ONE training step (forward and backward) pack every weight
    once.  The caller guarantees that no parameter changes inside the scope (train_step: the optimizer runs after it);
    outside a scope nothing is cached, so a parameter edited between two forwards can never meet a stale image."""
    _next = 1

    def __enter__(self):
        global _PACK_SCOPE
        self._prev = _PACK_SCOPE
        _PACK_SCOPE = weight_pack_cache._next
        weight_pack_cache._next += 1
        return self

    def __exit__(self, *exc):
        global _PACK_SCOPE
        _PACK_SCOPE = self._prev
        return False


def packed_weight(param, w, dgrad=False):
    """pack_conv_weight / pack_conv_weight_dgrad of ``w`` (the fp32 view of ``param``); cached on the parameter object
    for the lifetime of the enclosing ``weight_pack_cache`` scope."""
    if _PACK_SCOPE == 0:
        return pack_conv_weight_dgrad(w) if dgrad else pack_conv_weight(w)
    key = (_PACK_SCOPE, w.data_ptr(), tuple(w.shape))
    c = getattr(param, '_srb_packs', None)
    if c is None or c[0] != key:
        c = (key, {})
        param._srb_packs = c
    if dgrad not in c[1]:
        c[1][dgrad] = pack_conv_weight_dgrad(w) if dgrad else pack_conv_weight(w)
    return c[1][dgrad]


def pack_conv_weight(w):
    """fp32 [Cout,Cin,kh,kw] (or Linear [Cout,Cin]) CUDA tensor -> bf16 [taps][Cout][Cin]."""
    w = w.detach().float().contiguous()
    cout, cin = w.shape[0], w.shape[1]
    taps = w.numel() // (cout * cin)
    out = torch.empty(taps, cout, cin, dtype=torch.bfloat16, device=w.device)
    L.check(L.lib().srb_pack_conv_weight(L.ptr(w), L.ptr(out), cout, cin, taps, L.current_stream()), 'pack_conv_weight')
    return out


def conv_nhwc(x, cin, w_packed, bias=None, out=None, out32=None, act=L.ACT_NONE, slope=0.2, alpha1=1.0, res1=None,
              alpha2=1.0, res2=None, res2_cs=None, pixel_shuffle=False):
    """x: bf16 [B,H,W,Cpitch] (possibly a channel-offset view of a wider buffer); reads channels [0,cin).

    out / out32 / res1 / res2: NHWC tensors (views allowed; last-dim stride must be 1); pitch = stride(2).
    """
    B, H, W = x.shape[:3]
    taps, cout = w_packed.shape[0], w_packed.shape[1]
    assert w_packed.shape[2] == cin

    def pitch(t):
        return 0 if t is None else t.stride(2)

    def isf32(t):
        return int(t is not None and t.dtype == torch.float32)

    rc = L.lib().srb_conv_nhwc_bf16(
        L.ptr(x), x.stride(2), L.ptr(w_packed), L.ptr(bias), L.ptr(out), pitch(out), L.ptr(out32), pitch(out32),
        B, H, W, cin, cout, taps, act, slope, alpha1, L.ptr(res1), isf32(res1), pitch(res1), alpha2, L.ptr(res2),
        isf32(res2), pitch(res2), L.ptr(res2_cs), int(pixel_shuffle), L.current_stream())
    L.check(rc, 'conv_nhwc_bf16')


def conv_nhwc_train(x, cin, w_packed, bias=None, out=None, act=L.ACT_NONE, slope=0.2, res1=None, pre_out=None,
                    gelu_grad_of=None, alpha1_samples=None, res2=None, res2_cs=None):
    """conv_nhwc for the training path: optional ``pre_out`` (bf16, receives acc + bias before the activation) and
    ``gelu_grad_of`` (bf16 saved pre-activation: the result is multiplied by GELU' of it); ``alpha1_samples`` (fp32 [B]
    on the device): out = conv * alpha1_samples[image] + res1; ``res2`` (bf16) with ``res2_cs`` (fp32 [B, Cout]):
    out += res2 * res2_cs[image, channel].  Returns out."""
    B, H, W = x.shape[:3]
    taps, cout = w_packed.shape[0], w_packed.shape[1]
    assert w_packed.shape[2] == cin
    if out is None:
        out = torch.empty(B, H, W, cout, dtype=torch.bfloat16, device=x.device)
    pitch = lambda t: 0 if t is None else t.stride(2)
    L.check(L.lib().srb_conv_nhwc_bf16_train(
        L.ptr(x), x.stride(2), L.ptr(w_packed), L.ptr(bias), L.ptr(out), out.stride(2), B, H, W, cin, cout, taps, act,
        slope, L.ptr(res1), int(res1 is not None and res1.dtype == torch.float32), pitch(res1), L.ptr(pre_out),
        pitch(pre_out), L.ptr(gelu_grad_of), pitch(gelu_grad_of), L.ptr(alpha1_samples), L.ptr(res2), pitch(res2),
        L.ptr(res2_cs), L.current_stream()),
        'conv_nhwc_bf16_train')
    return out


def fold_ln_linear(w, b, gamma, beta):
    """LayerNorm affine folded into the following Linear: returns (bf16 [N,K] weight, fp32 [N] bias)."""
    w = w.detach().float().contiguous()
    n, k = w.shape
    wf = torch.empty(n, k, dtype=torch.bfloat16, device=w.device)
    bf = torch.empty(n, dtype=torch.float32, device=w.device)
    L.check(L.lib().srb_fold_ln_linear(L.ptr(w), L.ptr(b.float().contiguous()), L.ptr(gamma.float().contiguous()),
                                       L.ptr(beta.float().contiguous()), L.ptr(wf), L.ptr(bf), n, k,
                                       L.current_stream()), 'fold_ln_linear')
    return wf, bf


def mlp_branch(t, w1_folded, b1_folded, w2_packed, b2, cab=None, cs=None, out=None, tok_per_img=None, next_ln=None):
    """y = y0 + fc2(GELU(fc1(LN(y0)))), y0 = t + cab*cs[b].  t/cab/out: bf16 [ntok,96]; cs fp32 [B,96].
    next_ln=(gamma, beta): also returns xn = LayerNorm(y)*gamma + beta (bf16), the fused norm1 of the next block."""
    ntok = t.shape[0]
    if out is None:
        out = torch.empty_like(t)
    if tok_per_img is None:
        tok_per_img = ntok if cs is None else ntok // cs.shape[0]
    if next_ln is not None:
        xn = torch.empty_like(t)
        rc = L.lib().srb_mlp_branch_ln(L.ptr(t), L.ptr(cab), L.ptr(cs), L.ptr(w1_folded), L.ptr(b1_folded),
                                       L.ptr(w2_packed), L.ptr(b2), L.ptr(out), L.ptr(next_ln[0]), L.ptr(next_ln[1]),
                                       L.ptr(xn), ntok, tok_per_img, L.current_stream())
        L.check(rc, 'mlp_branch_ln')
        return out, xn
    rc = L.lib().srb_mlp_branch(L.ptr(t), L.ptr(cab), L.ptr(cs), L.ptr(w1_folded), L.ptr(b1_folded), L.ptr(w2_packed),
                                L.ptr(b2), L.ptr(out), ntok, tok_per_img, L.current_stream())
    L.check(rc, 'mlp_branch')
    return out


def expand_rel_bias(table, index, nk, scale=1.0):
    """table fp32 [T,6], index int64 [64,nk] -> bf16 bias in MMA fragment order [6*64*nk] (times `scale`)."""
    out = torch.empty(6 * 64 * nk, dtype=torch.bfloat16, device=table.device)
    L.check(L.lib().srb_expand_rel_bias(L.ptr(table.float().contiguous()), table.shape[0], L.ptr(index.contiguous()),
                                        nk, L.ptr(out), scale, L.current_stream()), 'expand_rel_bias')
    return out


def hab_attention(x, ln_g, ln_b, wqkv_packed, bqkv, bias_frag_log2e, wproj_packed, bproj, shift):
    """x: bf16 [B,H,W,96] -> (xn, t): LayerNorm(x) and x + proj(window attention(xn))."""
    B, H, W, _ = x.shape
    xn, t = torch.empty_like(x), torch.empty_like(x)
    L.check(L.lib().srb_hab_attention(L.ptr(x), L.ptr(ln_g), L.ptr(ln_b), L.ptr(wqkv_packed), L.ptr(bqkv),
                                      L.ptr(bias_frag_log2e), L.ptr(wproj_packed), L.ptr(bproj), L.ptr(xn), L.ptr(t),
                                      B, H, W, shift, L.current_stream()), 'hab_attention')
    return xn, t


def pack_conv_weight_dgrad(w):
    """fp32 [Cout,Cin,kh,kw] (or Linear [Cout,Cin]) -> bf16 [taps][Cin][Cout], taps reversed: weights of the
    data-gradient conv  dX = conv_nhwc(dY, Cout, pack_conv_weight_dgrad(W))."""
    w = w.detach().float().contiguous()
    cout, cin = w.shape[0], w.shape[1]
    taps = w.numel() // (cout * cin)
    out = torch.empty(taps, cin, cout, dtype=torch.bfloat16, device=w.device)
    L.check(L.lib().srb_pack_conv_weight_dgrad(L.ptr(w), L.ptr(out), cout, cin, taps, L.current_stream()),
            'pack_conv_weight_dgrad')
    return out


def conv_wgrad(x, cin, dy, cout, taps, dw=None, db=None, accumulate=False, want_bias=True):
    """x: bf16 [B,H,W,>=cin] layer input, dy: bf16 [B,H,W,>=cout] output gradient (channel-offset views allowed).
    Returns (dW fp32 [cout,cin,k,k] or [cout,cin], db fp32 [cout] or None)."""
    B, H, W = x.shape[:3]
    k = 3 if taps == 9 else 1
    if dw is None:
        assert not accumulate
        dw = torch.empty((cout, cin, k, k) if taps == 9 else (cout, cin), dtype=torch.float32, device=x.device)
    if db is None and want_bias:
        assert not accumulate
        db = torch.empty(cout, dtype=torch.float32, device=x.device)
    lib = L.lib()
    nbytes = lib.srb_conv_wgrad_scratch_bytes(B, H, W, cin, cout, taps)
    if nbytes == 0:
        raise RuntimeError('conv_wgrad: unsupported geometry (W must be a multiple of 32)')
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    L.check(lib.srb_conv_wgrad(L.ptr(x), x.stride(2), L.ptr(dy), dy.stride(2), L.ptr(dw), L.ptr(db), L.ptr(scratch),
                               B, H, W, cin, cout, taps, int(accumulate), L.current_stream()), 'conv_wgrad')
    return dw, db


def _npx(t):
    return t.shape[0] * t.shape[1] * t.shape[2]


def lrelu_bwd_(g, y, slope=0.2):
    """g *= (y > 0 ? 1 : slope) in place; g, y: bf16 NHWC (channel-offset views allowed), y = saved activation output."""
    L.check(L.lib().srb_lrelu_bwd(L.ptr(g), g.stride(2), L.ptr(y), y.stride(2), _npx(g), g.shape[3], slope,
                                  L.current_stream()), 'lrelu_bwd')
    return g


def axpby(out, a, x, b=0.0, y=None):
    """out = a*x + b*y on NHWC tensors / channel slices (bf16 or fp32 each); returns out."""
    f32 = lambda t: int(t is not None and t.dtype == torch.float32)
    L.check(L.lib().srb_axpby(L.ptr(out), f32(out), out.stride(2), float(a), L.ptr(x), f32(x), x.stride(2), float(b),
                              L.ptr(y), f32(y), 0 if y is None else y.stride(2), _npx(out), out.shape[3],
                              L.current_stream()), 'axpby')
    return out


def axpby_samples(out, a, x, b=0.0, y=None):
    """out[n] = a[n] * x[n] + b * y[n] on NHWC tensors; a: fp32 device tensor [B] (per-sample DropPath factors)."""
    f32 = lambda t: int(t is not None and t.dtype == torch.float32)
    assert a.dtype == torch.float32 and a.numel() == out.shape[0] and a.is_cuda
    L.check(L.lib().srb_axpby_samples(L.ptr(out), f32(out), out.stride(2), L.ptr(a), out.shape[1] * out.shape[2], L.ptr(x),
                                      f32(x), x.stride(2), float(b), L.ptr(y), f32(y), 0 if y is None else y.stride(2),
                                      _npx(out), out.shape[3], L.current_stream()), 'axpby_samples')
    return out


def upsample2x_bwd(g, out=None):
    """g: bf16 [B,2h,2w,C] -> bf16 [B,h,w,C], sum over each 2x2 block (backward of nearest 2x upsampling)."""
    B, H2, W2, C = g.shape
    if out is None:
        out = torch.empty(B, H2 // 2, W2 // 2, C, dtype=torch.bfloat16, device=g.device)
    L.check(L.lib().srb_upsample2x_bwd(L.ptr(g), g.stride(2), L.ptr(out), out.stride(2), B, H2 // 2, W2 // 2, C,
                                       L.current_stream()), 'upsample2x_bwd')
    return out


def upsample2x_nearest(x, out=None):
    """bf16 [B,h,w,64] (channel-offset view allowed) -> dense bf16 [B,2h,2w,64], nearest neighbour."""
    B, h, w, C = x.shape
    assert C == 64
    if out is None:
        out = torch.empty(B, 2 * h, 2 * w, 64, dtype=torch.bfloat16, device=x.device)
    L.check(L.lib().srb_upsample2x_nearest(L.ptr(x), x.stride(2), L.ptr(out), B, h, w, L.current_stream()),
            'upsample2x_nearest')
    return out


def conv3x3_cout1(x, w, bias):
    """bf16 [B,H,W,64] -> fp32 [B,H,W]; w fp32 [1,64,3,3], bias fp32 [1]."""
    B, H, W, _ = x.shape
    out = torch.empty(B, H, W, dtype=torch.float32, device=x.device)
    L.check(L.lib().srb_conv3x3_cout1(L.ptr(x), x.stride(2), L.ptr(w.detach().float().contiguous()),
                                      L.ptr(bias.detach().float().contiguous()), L.ptr(out), B, H, W,
                                      L.current_stream()), 'conv3x3_cout1')
    return out


def layernorm96(x, gamma, beta):
    """bf16 [ntok,96] -> bf16 [ntok,96], LayerNorm over the 96 channels (eps 1e-5)."""
    y = torch.empty_like(x)
    L.check(L.lib().srb_layernorm96(L.ptr(x), L.ptr(gamma), L.ptr(beta), L.ptr(y), x.shape[0], L.current_stream()),
            'layernorm96')
    return y


def layernorm96_bwd(x, g, gamma, res=None):
    """x, g: bf16 [ntok,96] (LN input, gradient of the LN output).  Returns (dx bf16 (+ res), dgamma fp32, dbeta fp32)."""
    dx = torch.empty_like(x)
    dg = torch.empty(96, dtype=torch.float32, device=x.device)
    db = torch.empty(96, dtype=torch.float32, device=x.device)
    lib = L.lib()
    scratch = torch.empty(lib.srb_layernorm96_bwd_scratch_bytes(), dtype=torch.uint8, device=x.device)
    L.check(lib.srb_layernorm96_bwd(L.ptr(x), L.ptr(g), L.ptr(gamma), L.ptr(res),
                                    int(res is not None and res.dtype == torch.float32), L.ptr(dx), L.ptr(dg), L.ptr(db),
                                    L.ptr(scratch), x.shape[0], 0, L.current_stream()), 'layernorm96_bwd')
    return dx, dg, db


def gelu(pre, grad=None, out=None):
    """pre: bf16 [...,C] (last-dim-contiguous 2-D / NHWC slice).  grad None: GELU(pre); else grad * GELU'(pre)."""
    p2 = pre.reshape(-1, pre.shape[-1]) if pre.is_contiguous() else pre
    npx = pre.numel() // pre.shape[-1]
    if out is None:
        out = torch.empty(pre.shape, dtype=torch.bfloat16, device=pre.device)
    pitch = lambda t: t.stride(-2)
    L.check(L.lib().srb_gelu(L.ptr(out), pitch(out), L.ptr(pre), pitch(pre), L.ptr(grad),
                             0 if grad is None else pitch(grad), npx, pre.shape[-1], L.current_stream()), 'gelu')
    return out


def gather_rel_bias(table, index):
    """relative_position_bias_table fp32 [225,6] + relative_position_index int64 [64,64] -> fp32 [6,64,64] (a gather)."""
    return table.detach().float()[index.view(-1)].view(64, 64, 6).permute(2, 0, 1).contiguous()


def window_attention(qkv, bias, shift):
    """qkv bf16 [B,H,W,288], bias fp32 [6,64,64] -> bf16 [B,H,W,96] (attention output before proj)."""
    B, H, W, _ = qkv.shape
    out = torch.empty(B, H, W, 96, dtype=torch.bfloat16, device=qkv.device)
    L.check(L.lib().srb_window_attention(L.ptr(qkv), L.ptr(bias), L.ptr(out), B, H, W, shift, L.current_stream()),
            'window_attention')
    return out


def window_attention_bwd(qkv, bias, index, dout, shift, dtable=None):
    """Returns (dqkv bf16 [B,H,W,288], dtable fp32 [225,6]); dtable is accumulated into when given."""
    B, H, W, _ = qkv.shape
    dqkv = torch.empty_like(qkv)
    if dtable is None:
        dtable = torch.zeros(225, 6, dtype=torch.float32, device=qkv.device)
    idx = index.to(torch.int32).contiguous()
    L.check(L.lib().srb_window_attention_bwd(L.ptr(qkv), L.ptr(bias), L.ptr(idx), L.ptr(dout), L.ptr(dqkv), L.ptr(dtable),
                                             B, H, W, shift, L.current_stream()), 'window_attention_bwd')
    return dqkv, dtable


def gather_rel_bias_oca(table, index):
    """OCAB: table fp32 [361,6] + index int64 [64,144] -> fp32 [6,64,144]."""
    return table.detach().float()[index.view(-1)].view(64, 144, 6).permute(2, 0, 1).contiguous()


def overlap_attention(qkv, bias):
    """qkv bf16 [B,H,W,288], bias fp32 [6,64,144] -> bf16 [B,H,W,96] (OCAB attention output before proj)."""
    B, H, W, _ = qkv.shape
    out = torch.empty(B, H, W, 96, dtype=torch.bfloat16, device=qkv.device)
    L.check(L.lib().srb_overlap_attention(L.ptr(qkv), L.ptr(bias), L.ptr(out), B, H, W, L.current_stream()),
            'overlap_attention')
    return out


def overlap_attention_infer(qkv_qscaled, bias_frag_log2e):
    """The OCAB core of the inference forward (csrc/attn.cu).  qkv bf16 [B,H,W,288] whose q part is already scaled by
    0.25 * log2(e); bias_frag_log2e = expand_rel_bias(table, index, 144, scale=log2(e)).  -> bf16 [B,H,W,96]."""
    B, H, W, _ = qkv_qscaled.shape
    out = torch.empty(B, H, W, 96, dtype=torch.bfloat16, device=qkv_qscaled.device)
    L.check(L.lib().srb_overlap_attention_infer(L.ptr(qkv_qscaled), L.ptr(bias_frag_log2e), L.ptr(out), B, H, W,
                                                L.current_stream()), 'overlap_attention_infer')
    return out


def overlap_attention_bwd(qkv, bias, index, dout, dtable=None):
    """Returns (dqkv bf16 [B,H,W,288], dtable fp32 [361,6]); dtable is accumulated into when given."""
    B, H, W, _ = qkv.shape
    dqkv = torch.empty_like(qkv)
    if dtable is None:
        dtable = torch.zeros(361, 6, dtype=torch.float32, device=qkv.device)
    # HAT.calculate_rpi_oca yields indices in [-200, 160]: negative ones address the table from its end (Python indexing)
    idx = (index % 361).to(torch.int32).contiguous()
    lib = L.lib()
    scratch = torch.empty(lib.srb_overlap_attention_bwd_scratch_bytes(B, H, W), dtype=torch.uint8, device=qkv.device)
    L.check(lib.srb_overlap_attention_bwd(L.ptr(qkv), L.ptr(bias), L.ptr(idx), L.ptr(dout), L.ptr(dqkv), L.ptr(dtable),
                                          L.ptr(scratch), B, H, W, L.current_stream()), 'overlap_attention_bwd')
    return dqkv, dtable


def channel_attention(y, w1, b1, w2, b2, want_out=True):
    """y bf16 [B,HW,96]; w1 [3,96(,1,1)], w2 [96,3(,1,1)] fp32.  Returns (out bf16, saved = (pooled, hid, scale));
    want_out=False: out is None and only the statistics are computed (scale [B,96] is applied by the caller)."""
    B, HW, _ = y.shape
    dev = y.device
    out = torch.empty_like(y) if want_out else None
    pooled = torch.empty(B, 96, dtype=torch.float32, device=dev)
    hid = torch.empty(B, 3, dtype=torch.float32, device=dev)
    scale = torch.empty(B, 96, dtype=torch.float32, device=dev)
    lib = L.lib()
    scratch = torch.empty(lib.srb_channel_attention_scratch_bytes(B, HW), dtype=torch.uint8, device=dev)
    f = lambda t: t.detach().float().contiguous()
    L.check(lib.srb_channel_attention(L.ptr(y), L.ptr(f(w1)), L.ptr(f(b1)), L.ptr(f(w2)), L.ptr(f(b2)), L.ptr(out),
                                      L.ptr(pooled), L.ptr(hid), L.ptr(scale), L.ptr(scratch), B, HW,
                                      L.current_stream()), 'channel_attention')
    return out, (pooled, hid, scale)


def channel_attention_bwd(y, g, w1, w2, saved, g_scale=1.0):
    """Returns (dy bf16 [B,HW,96], dw1 [3,96], db1 [3], dw2 [96,3], db2 [96]) for the incoming gradient g_scale * g."""
    B, HW, _ = y.shape
    dev = y.device
    pooled, hid, scale = saved
    dy = torch.empty_like(y)
    mk = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
    dw1, db1, dw2, db2, dpooled = mk(3, 96), mk(3), mk(96, 3), mk(96), mk(B, 96)
    lib = L.lib()
    scratch = torch.empty(lib.srb_channel_attention_scratch_bytes(B, HW), dtype=torch.uint8, device=dev)
    f = lambda t: t.detach().float().contiguous()
    L.check(lib.srb_channel_attention_bwd(L.ptr(y), L.ptr(g), L.ptr(f(w1)), L.ptr(f(w2)), L.ptr(pooled), L.ptr(hid),
                                          L.ptr(scale), L.ptr(dy), L.ptr(dw1), L.ptr(db1), L.ptr(dw2), L.ptr(db2),
                                          L.ptr(dpooled), L.ptr(scratch), B, HW, 0, float(g_scale), L.current_stream()),
            'channel_attention_bwd')
    return dy, dw1, db1, dw2, db2


def blur(x, mode, out_dtype=torch.float32):
    """AntiCheckerboardLayer.forward (reference src/architecture.py:26-27) on fp32 [B,1,H,W] / [B,H,W]; mode 'light' |
    'balanced' | 'strong'.  The zero-padded binomial operator is symmetric: its backward is the same call."""
    x = x.contiguous()
    if x.dtype != torch.float32:
        raise TypeError('blur: fp32 input expected')
    H, W = x.shape[-2:]
    out = torch.empty(x.shape, dtype=out_dtype, device=x.device)
    L.check(L.lib().srb_blur(L.ptr(x), L.ptr(out), L.SRB_U16 if out_dtype == torch.uint16 else L.SRB_F32,
                             L.SMOOTHING[mode], x.numel() // (H * W), H, W, L.current_stream()), 'blur')
    return out


_BICUBIC_TABLES = {}


def _cubic_taps(n_in, n_out):
    """tap tables of ATen's upsample_bicubic2d (align_corners=False, A=-0.75, no scale_factor): source coordinate
    (dst + 0.5) * n_in / n_out - 0.5 (not clamped), 4 taps at floor - 1 .. floor + 2 clamped to the image."""
    A = -0.75
    scale = torch.tensor(n_in / n_out, dtype=torch.float32)
    real = scale * (torch.arange(n_out, dtype=torch.float32) + 0.5) - 0.5
    fl = torch.floor(real)
    t = real - fl
    c1 = lambda v: ((A + 2) * v - (A + 3)) * v * v + 1
    c2 = lambda v: ((A * v - 5 * A) * v + 8 * A) * v - 4 * A
    w = torch.stack([c2(t + 1.0), c1(t), c1(1.0 - t), c2((1.0 - t) + 1.0)], 1).float()
    idx = (fl.long()[:, None] + torch.arange(-1, 3)[None]).clamp_(0, n_in - 1).int()
    return idx, w


def _cubic_taps_aa(n_in, n_out):
    """tap tables of ATen's anti-aliased bicubic (F.interpolate(..., mode='bicubic', antialias=True), the call of the
    stale Backup/src/architecture.py:91-92): Keys filter with A = -0.5 stretched by max(scale, 1), taps
    [int(center - support + 0.5), int(center + support + 0.5)) clipped to the image and renormalised."""
    A = -0.5
    scale = n_in / n_out
    support = 2.0 * max(scale, 1.0)
    inv = 1.0 / max(scale, 1.0)
    K = int(math.ceil(support)) * 2 + 1

    def filt(x):
        x = abs(x)
        if x < 1.0:
            return ((A + 2.0) * x - (A + 3.0)) * x * x + 1.0
        if x < 2.0:
            return (((x - 5.0) * x + 8.0) * x - 4.0) * A
        return 0.0

    idx = torch.zeros(n_out, K, dtype=torch.int32)
    w = torch.zeros(n_out, K, dtype=torch.float32)
    for o in range(n_out):
        center = scale * (o + 0.5)
        xmin = max(0, int(center - support + 0.5))
        xsize = min(n_in, int(center + support + 0.5)) - xmin
        ws = [filt((j + xmin - center + 0.5) * inv) for j in range(xsize)]
        tot = sum(ws)
        for j in range(xsize):
            idx[o, j] = xmin + j
            w[o, j] = ws[j] / tot
        idx[o, xsize:] = xmin  # padding taps: weight 0 on a valid index
    return idx, w


def _transpose_taps(idx, w, n_in):
    """tables of the transposed operator: for every input position, the outputs it feeds and their weights (padded)"""
    n_out, K = idx.shape
    lists = [[] for _ in range(n_in)]
    for o in range(n_out):
        for k in range(K):
            lists[int(idx[o, k])].append((o, float(w[o, k])))
    Kt = max(len(l) for l in lists)
    it = torch.zeros(n_in, Kt, dtype=torch.int32)
    wt = torch.zeros(n_in, Kt, dtype=torch.float32)
    for i, l in enumerate(lists):
        for k, (o, v) in enumerate(l):
            it[i, k], wt[i, k] = o, v
    return it, wt


def bicubic_tables(n_in, n_out, device, transpose=False, antialias=False):
    key = (n_in, n_out, str(device), transpose, antialias)
    if key not in _BICUBIC_TABLES:
        idx, w = _cubic_taps_aa(n_in, n_out) if antialias else _cubic_taps(n_in, n_out)
        if transpose:
            idx, w = _transpose_taps(idx, w, n_in)
            if idx.shape[1] > 64:
                raise ValueError(f'bicubic backward {n_out} -> {n_in}: more than 64 outputs per input pixel')
        _BICUBIC_TABLES[key] = (idx.contiguous().to(device), w.contiguous().to(device))
    return _BICUBIC_TABLES[key]


def resample_separable(x, Ho, Wo, ty, tx):
    """fp32 [B,1,Hi,Wi] -> [B,1,Ho,Wo] through per-axis (index, weight) tap tables (see bicubic_tables)."""
    x = x.contiguous()
    B, Hi, Wi = x.numel() // (x.shape[-2] * x.shape[-1]), x.shape[-2], x.shape[-1]
    out = torch.empty(*x.shape[:-2], Ho, Wo, dtype=torch.float32, device=x.device)
    K = ty[0].shape[1]
    assert tx[0].shape[1] == K and ty[0].shape[0] == Ho and tx[0].shape[0] == Wo
    L.check(L.lib().srb_resample_separable(L.ptr(x), L.ptr(out), B, Hi, Wi, Ho, Wo, L.ptr(ty[0]), L.ptr(ty[1]),
                                           L.ptr(tx[0]), L.ptr(tx[1]), K, L.current_stream()), 'resample_separable')
    return out


def bicubic_resize(x, Ho, Wo, antialias=False):
    """F.interpolate(x, size=(Ho, Wo), mode='bicubic', align_corners=False) on fp32 [B,1,H,W] (src/architecture.py:80-81);
    ``antialias=True`` = the stale variant's call (Backup/src/architecture.py:91-92)."""
    Hi, Wi = x.shape[-2:]
    ty, tx = bicubic_tables(Hi, Ho, x.device, False, antialias), bicubic_tables(Wi, Wo, x.device, False, antialias)
    K = max(ty[0].shape[1], tx[0].shape[1])
    pad = lambda t: tuple(torch.nn.functional.pad(a, (0, K - a.shape[1])).contiguous() for a in t)
    return resample_separable(x, Ho, Wo, pad(ty), pad(tx))


def bicubic_resize_bwd(g, Hi, Wi):
    """gradient of bicubic_resize w.r.t. its input: the transposed separable operator applied to g [B,1,Ho,Wo]."""
    Ho, Wo = g.shape[-2:]
    ty, tx = bicubic_tables(Hi, Ho, g.device, True), bicubic_tables(Wi, Wo, g.device, True)
    K = max(ty[0].shape[1], tx[0].shape[1])
    pad = lambda t: tuple(torch.nn.functional.pad(a, (0, K - a.shape[1])).contiguous() for a in t)
    return resample_separable(g, Hi, Wi, pad(ty), pad(tx))
