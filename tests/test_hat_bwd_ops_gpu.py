"""HAT (stage 2) backward building blocks against torch autograd in fp32 on the same bf16-rounded inputs:
LayerNorm(96) forward / backward (+ fused shortcut gradient), exact-erf GELU forward / backward.
(the operators the stage-2 training path of training_hat.py is assembled from.)"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


# >= 16384 tokens with a bf16 / no residual take the bulk-copy pipeline kernel of the backward (ragged last tile included)
@pytest.mark.parametrize('ntok,with_res', [(64, None), (4096 + 37, 'bf16'), (148 * 4 * 16 * 3 + 5, 'f32'),
                                           (16384 + 128 * 37 + 19, 'bf16'), (300 * 128 * 3 + 77, None)])
def test_layernorm96_forward_backward(cuda, ntok, with_res):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(ntok)
    x = (torch.randn(ntok, 96, generator=g) * 1.3 + 0.2).to(cuda).bfloat16()
    gy = torch.randn(ntok, 96, generator=g).to(cuda).bfloat16()
    gamma = (1 + 0.3 * torch.randn(96, generator=g)).to(cuda)
    beta = (0.2 * torch.randn(96, generator=g)).to(cuda)
    res = None
    if with_res:
        res = torch.randn(ntok, 96, generator=g).to(cuda)
        res = res.bfloat16() if with_res == 'bf16' else res
    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y_ref = F.layer_norm(xr, (96,), gr, br, 1e-5)
    y_ref.backward(gy.float())
    y = ops.layernorm96(x, gamma, beta)
    err = (y.float() - y_ref.detach()).abs()  # bf16 output: half an ulp is 2^-9 relative
    assert (err <= 4e-3 * y_ref.detach().abs() + 1e-2).all(), err.max().item()
    dx, dg, db = ops.layernorm96_bwd(x, gy, gamma, res=res)
    want = xr.grad + (res.float() if res is not None else 0)
    err = (dx.float() - want).abs()
    assert (err <= 1e-2 * want.abs() + 2e-2).all(), err.max().item()
    scale = ntok ** 0.5
    assert (dg - gr.grad).abs().max().item() <= 2e-3 * scale + 1e-3
    assert (db - br.grad).abs().max().item() <= 2e-3 * scale + 1e-3
    dx2, dg2, db2 = ops.layernorm96_bwd(x, gy, gamma, res=res)
    assert torch.equal(dg, dg2) and torch.equal(db, db2) and torch.equal(dx, dx2)  # deterministic


def test_gelu_forward_backward(cuda):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(3)
    buf = (torch.randn(2, 16, 32, 96, generator=g) * 2).to(cuda).bfloat16()
    pre = buf[..., 32:64]  # a channel slice of a wider NHWC buffer
    gy = torch.randn(2, 16, 32, 32, generator=g).to(cuda).bfloat16()
    pr = pre.float().requires_grad_(True)
    y_ref = F.gelu(pr)
    y_ref.backward(gy.float())
    y = ops.gelu(pre)
    assert (y.float() - y_ref.detach()).abs().max().item() <= 2e-2
    d = ops.gelu(pre, grad=gy)
    err = (d.float() - pr.grad).abs()
    assert (err <= 1e-2 * pr.grad.abs() + 1e-2).all(), err.max().item()
    ops.gelu(pre, grad=gy, out=gy)  # in place on the gradient
    assert torch.equal(gy, d)


def _ref_window_attention(qkv, table, index, shift):
    """torch restatement of HAB's attention core (upstream hat_arch.py: roll, window_partition, WindowAttention without the
    qkv / proj linears, window_reverse, roll back) on fp32 tensors; differentiable."""
    B, H, W, _ = qkv.shape
    x = torch.roll(qkv, (-shift, -shift), (1, 2)) if shift else qkv
    xw = x.view(B, H // 8, 8, W // 8, 8, 288).permute(0, 1, 3, 2, 4, 5).reshape(-1, 64, 3, 6, 16).permute(2, 0, 3, 1, 4)
    q, k, v = xw[0] * 0.25, xw[1], xw[2]                       # [nW, 6, 64, 16]
    attn = q @ k.transpose(-2, -1) + table[index.view(-1)].view(64, 64, 6).permute(2, 0, 1).unsqueeze(0)
    if shift:
        img = torch.zeros(H, W, device=qkv.device)
        cnt = 0
        for hs in (slice(0, -8), slice(-8, -4), slice(-4, None)):
            for ws in (slice(0, -8), slice(-8, -4), slice(-4, None)):
                img[hs, ws] = cnt
                cnt += 1
        mw = img.view(H // 8, 8, W // 8, 8).permute(0, 2, 1, 3).reshape(-1, 64)
        mask = (mw.unsqueeze(1) - mw.unsqueeze(2)).ne(0).float() * -100.0   # [nW_img, 64, 64]
        attn = attn.view(B, -1, 6, 64, 64) + mask.view(1, -1, 1, 64, 64)
        attn = attn.view(-1, 6, 64, 64)
    o = (attn.softmax(-1) @ v).transpose(1, 2).reshape(-1, 8, 8, 96)
    o = o.view(B, H // 8, W // 8, 8, 8, 96).permute(0, 1, 3, 2, 4, 5).reshape(B, H, W, 96)
    return torch.roll(o, (shift, shift), (1, 2)) if shift else o


@pytest.mark.parametrize('B,H,W,shift', [(1, 16, 16, 0), (2, 16, 32, 4), (1, 32, 24, 4), (3, 24, 16, 0)])
def test_window_attention_forward_backward(cuda, B, H, W, shift):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(B * 1000 + H + shift)
    qkv = (torch.randn(B, H, W, 288, generator=g) * 1.2).to(cuda).bfloat16()
    table = (torch.randn(225, 6, generator=g) * 0.5).to(cuda)
    coords = torch.stack(torch.meshgrid(torch.arange(8), torch.arange(8), indexing='ij')).flatten(1)
    rel = (coords[:, :, None] - coords[:, None, :]).permute(1, 2, 0) + 7
    index = (rel[..., 0] * 15 + rel[..., 1]).to(cuda)          # relative_position_index of an 8x8 window
    dout = torch.randn(B, H, W, 96, generator=g).to(cuda).bfloat16()
    qr = qkv.float().requires_grad_(True)
    tr = table.clone().requires_grad_(True)
    ref = _ref_window_attention(qr, tr, index, shift)
    ref.backward(dout.float())
    bias = ops.gather_rel_bias(table, index)
    out = ops.window_attention(qkv, bias, shift)
    err = (out.float() - ref.detach()).abs()
    assert (err <= 1e-2 * ref.detach().abs() + 2e-2).all(), err.max().item()
    dqkv, dtab = ops.window_attention_bwd(qkv, bias, index, dout, shift)
    err = (dqkv.float() - qr.grad).abs()
    assert (err <= 2e-2 * qr.grad.abs() + 3e-2).all(), err.max().item()
    assert (dtab - tr.grad).abs().max().item() <= 2e-2 * tr.grad.abs().max().item() + 1e-2


@pytest.mark.parametrize('B,HW', [(1, 256), (3, 32 * 40), (2, 64 * 64 + 64)])
def test_channel_attention_forward_backward(cuda, B, HW):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(B * 7 + HW)
    y = (torch.randn(B, HW, 96, generator=g) + 0.3).to(cuda).bfloat16()
    gy = torch.randn(B, HW, 96, generator=g).to(cuda).bfloat16()
    w1 = (torch.randn(3, 96, generator=g) * 0.3).to(cuda)
    b1 = (torch.randn(3, generator=g) * 0.3 + 0.2).to(cuda)
    w2 = (torch.randn(96, 3, generator=g) * 0.5).to(cuda)
    b2 = (torch.randn(96, generator=g) * 0.3).to(cuda)
    yr = y.float().requires_grad_(True)
    P = [t.clone().requires_grad_(True) for t in (w1, b1, w2, b2)]
    pooled = yr.mean(1)
    sc = torch.sigmoid(torch.relu(pooled @ P[0].t() + P[1]) @ P[2].t() + P[3])
    ref = yr * sc.unsqueeze(1)
    ref.backward(gy.float())
    out, saved = ops.channel_attention(y, w1, b1, w2, b2)
    err = (out.float() - ref.detach()).abs()
    assert (err <= 1e-2 * ref.detach().abs() + 1e-2).all(), err.max().item()
    for sc_g in (1.0, 0.25):  # g_scale: the gradient of sc_g * out without a scaling pass over gy (HAB's conv_scale)
        dy, dw1, db1, dw2, db2 = ops.channel_attention_bwd(y, gy, w1, w2, saved, g_scale=sc_g)
        want_dy = sc_g * yr.grad
        err = (dy.float() - want_dy).abs()
        assert (err <= 1e-2 * want_dy.abs() + 1e-2 * sc_g).all(), err.max().item()
        for got, want in ((dw1, P[0].grad), (db1, P[1].grad), (dw2, P[2].grad), (db2, P[3].grad)):
            want = sc_g * want
            assert (got - want).abs().max().item() <= 1e-2 * want.abs().max().item() + 1e-3 * sc_g, (got - want).abs().max().item()


def _ref_overlap_attention(qkv, table, index):
    """OCAB.forward's attention core (oracle/hat.py OCAB.forward) on a [B,H,W,288] qkv tensor, fp32."""
    B, H, W, _ = qkv.shape
    c, ws, ows = 96, 8, 12
    q = qkv[..., :96]
    kv = qkv[..., 96:].permute(0, 3, 1, 2)  # b, 2c, h, w
    qw = q.view(B, H // ws, ws, W // ws, ws, c).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, c)
    kvw = torch.nn.functional.unfold(kv, kernel_size=ows, stride=ws, padding=2)
    nw = kvw.shape[-1]
    kvw = kvw.view(B, 2, c, ows, ows, nw).permute(1, 0, 5, 3, 4, 2).reshape(2, B * nw, ows * ows, c)
    b_ = qw.shape[0]
    qh = qw.reshape(b_, 64, 6, 16).permute(0, 2, 1, 3) * 0.25
    kh = kvw[0].reshape(b_, 144, 6, 16).permute(0, 2, 1, 3)
    vh = kvw[1].reshape(b_, 144, 6, 16).permute(0, 2, 1, 3)
    bias = table[index.view(-1)].view(64, 144, 6).permute(2, 0, 1)
    attn = torch.softmax(qh @ kh.transpose(-2, -1) + bias.unsqueeze(0), dim=-1)
    o = (attn @ vh).transpose(1, 2).reshape(b_, 64, c)
    return o.view(B, H // ws, W // ws, ws, ws, c).permute(0, 1, 3, 2, 4, 5).reshape(B, H, W, c)


@pytest.mark.parametrize('B,H,W', [(1, 8, 8), (1, 16, 16), (2, 16, 32), (3, 24, 16)])
def test_overlap_attention_forward_backward(cuda, B, H, W):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(B * 1000 + H + W)
    qkv = (torch.randn(B, H, W, 288, generator=g) * 1.2).to(cuda).bfloat16()
    table = (torch.randn(361, 6, generator=g) * 0.5).to(cuda)
    co = torch.stack(torch.meshgrid(torch.arange(8), torch.arange(8), indexing='ij')).flatten(1)
    ce = torch.stack(torch.meshgrid(torch.arange(12), torch.arange(12), indexing='ij')).flatten(1)
    rel = (ce[:, None, :] - co[:, :, None]).permute(1, 2, 0) + (8 - 12 + 1)   # HAT.calculate_rpi_oca
    index = (rel[..., 0] * 19 + rel[..., 1]).to(cuda)
    dout = torch.randn(B, H, W, 96, generator=g).to(cuda).bfloat16()
    qr = qkv.float().requires_grad_(True)
    tr = table.clone().requires_grad_(True)
    ref = _ref_overlap_attention(qr, tr, index)
    ref.backward(dout.float())
    bias = ops.gather_rel_bias_oca(table, index)
    out = ops.overlap_attention(qkv, bias)
    err = (out.float() - ref.detach()).abs()
    assert (err <= 1e-2 * ref.detach().abs() + 2e-2).all(), err.max().item()
    dqkv, dtab = ops.overlap_attention_bwd(qkv, bias, index, dout)
    err = (dqkv.float() - qr.grad).abs()
    assert (err <= 2e-2 * qr.grad.abs() + 4e-2).all(), err.max().item()
    assert (dtab - tr.grad).abs().max().item() <= 2e-2 * tr.grad.abs().max().item() + 1e-2


@pytest.mark.parametrize('B,H,W', [(1, 8, 8), (1, 16, 16), (2, 16, 32), (3, 24, 16), (1, 64, 64)])
def test_overlap_attention_inference_kernel(cuda, B, H, W):
    """The OCAB core every inference forward launches (csrc/attn.cu, C ABI srb_overlap_attention_infer) against the same
    fp32 restatement of upstream OCAB.forward, with a sigma-0.5 bias table: border windows included (8x8 is all border),
    so the zero-padded keys that stay in the softmax with logit = bias carry real weight."""
    import math
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(7 * B * 1000 + H + W)
    qkv = (torch.randn(B, H, W, 288, generator=g) * 1.2).to(cuda).bfloat16()
    table = (torch.randn(361, 6, generator=g) * 0.5).to(cuda)
    co = torch.stack(torch.meshgrid(torch.arange(8), torch.arange(8), indexing='ij')).flatten(1)
    ce = torch.stack(torch.meshgrid(torch.arange(12), torch.arange(12), indexing='ij')).flatten(1)
    rel = (ce[:, None, :] - co[:, :, None]).permute(1, 2, 0) + (8 - 12 + 1)
    index = (rel[..., 0] * 19 + rel[..., 1]).to(cuda)
    ref = _ref_overlap_attention(qkv.float(), table, index)
    # the inference schedule folds head_dim^-0.5 * log2(e) into the q rows of the qkv linear: emulate on the bf16 tensor
    qs = qkv.clone()
    qs[..., :96] = (qkv[..., :96].float() * (0.25 * math.log2(math.e))).bfloat16()
    bias = ops.expand_rel_bias(table, index, 144, scale=math.log2(math.e))
    out = ops.overlap_attention_infer(qs, bias)
    torch.cuda.synchronize()
    err = (out.float() - ref).abs()
    bad = (err > 2e-2 * ref.abs() + 3e-2).sum().item()
    assert bad == 0, f'{bad} mismatches, max err {err.max().item():.4g}'
    assert err.mean().item() < 6e-3, err.mean().item()
