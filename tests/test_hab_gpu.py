"""Parity of the fused HAB attention branch (C ABI srb_hab_attention: LayerNorm -> qkv -> (shifted) window attention
with relative-position bias and shift mask -> proj -> + shortcut) against a plain PyTorch fp32 restatement of the
upstream HAT code path (hat_arch.py HAB / WindowAttention / calculate_mask; reached from src/architecture.py:78).
Tolerance: same bf16-rounded inputs and weights; the kernel rounds xn, q/k/v, P and the attention output to bf16."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
WS, HEADS, C = 8, 6, 96


def _rpi():
    ch, cw = torch.arange(WS), torch.arange(WS)
    coords = torch.stack(torch.meshgrid(ch, cw, indexing='ij')).flatten(1)
    rel = (coords[:, :, None] - coords[:, None, :]).permute(1, 2, 0).contiguous()
    rel[:, :, 0] += WS - 1
    rel[:, :, 1] += WS - 1
    rel[:, :, 0] *= 2 * WS - 1
    return rel.sum(-1)  # [64, 64]


def _partition(x):  # [B,H,W,C] -> [nW*B, 64, C]
    B, H, W, Cc = x.shape
    x = x.view(B, H // WS, WS, W // WS, WS, Cc)
    return x.permute(0, 1, 3, 2, 4, 5).contiguous().view(-1, WS * WS, Cc)


def _reverse(w, B, H, W):
    x = w.view(B, H // WS, W // WS, WS, WS, -1)
    return x.permute(0, 1, 3, 2, 4, 5).contiguous().view(B, H, W, -1)


def _mask(H, W, shift, dev):
    img = torch.zeros(1, H, W, 1, device=dev)
    cnt = 0
    for hs in (slice(0, -WS), slice(-WS, -shift), slice(-shift, None)):
        for ws_ in (slice(0, -WS), slice(-WS, -shift), slice(-shift, None)):
            img[:, hs, ws_, :] = cnt
            cnt += 1
    mw = _partition(img).squeeze(-1)
    am = mw.unsqueeze(1) - mw.unsqueeze(2)
    return am.masked_fill(am != 0, -100.0).masked_fill(am == 0, 0.0)  # [nW, 64, 64]


def _ref(x, g, b, wqkv, bqkv, table, rpi, wproj, bproj, shift):
    B, H, W, _ = x.shape
    x = x.float()
    xn = F.layer_norm(x, (C,), g, b, 1e-5)
    xs = torch.roll(xn, (-shift, -shift), (1, 2)) if shift else xn
    xw = _partition(xs)
    qkv = (xw @ wqkv.t() + bqkv).view(-1, 64, 3, HEADS, C // HEADS).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0] * (C // HEADS) ** -0.5, qkv[1], qkv[2]
    attn = q @ k.transpose(-2, -1) + table[rpi.view(-1)].view(64, 64, HEADS).permute(2, 0, 1).unsqueeze(0)
    if shift:
        m = _mask(H, W, shift, x.device)
        nW = m.shape[0]
        attn = (attn.view(-1, nW, HEADS, 64, 64) + m.unsqueeze(1).unsqueeze(0)).view(-1, HEADS, 64, 64)
    o = (attn.softmax(-1) @ v).transpose(1, 2).reshape(-1, 64, C) @ wproj.t() + bproj
    o = _reverse(o, B, H, W)
    if shift:
        o = torch.roll(o, (shift, shift), (1, 2))
    return xn, x + o


# the last two geometries give every persistent CTA several tiles (the cross-tile pipelining of the kernel) and an odd
# number of windows (the last tile's second window is empty)
@pytest.mark.parametrize('B,H,W,shift', [(1, 16, 16, 0), (1, 16, 16, 4), (2, 32, 64, 4), (3, 64, 32, 0), (2, 64, 64, 4),
                                         (3, 256, 256, 4), (1, 24, 24, 4), (5, 72, 88, 0)])
def test_hab_attention(cuda, B, H, W, shift):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(B * 1000 + H + W + shift)
    bq = lambda t: t.to(torch.bfloat16).float()  # weights the kernel sees
    x = (torch.randn(B, H, W, C, generator=g) * 1.2 + 0.2).to(cuda).to(torch.bfloat16)
    ln_g = (1 + 0.2 * torch.randn(C, generator=g)).to(cuda)
    ln_b = (0.1 * torch.randn(C, generator=g)).to(cuda)
    wqkv = bq(torch.randn(3 * C, C, generator=g) / C ** 0.5).to(cuda)
    bqkv = (0.2 * torch.randn(3 * C, generator=g)).to(cuda)
    wproj = bq(torch.randn(C, C, generator=g) / C ** 0.5).to(cuda)
    bproj = (0.2 * torch.randn(C, generator=g)).to(cuda)
    table = (0.5 * torch.randn((2 * WS - 1) ** 2, HEADS, generator=g)).to(cuda)
    rpi = _rpi().to(cuda)
    xn_ref, t_ref = _ref(x, ln_g, ln_b, wqkv, bqkv, table, rpi, wproj, bproj, shift)
    bias = ops.expand_rel_bias(table, rpi, 64, scale=math.log2(math.e))
    xn, t = ops.hab_attention(x, ln_g, ln_b, ops.pack_conv_weight(wqkv), bqkv, bias, ops.pack_conv_weight(wproj), bproj,
                              shift)
    torch.cuda.synchronize()
    e_xn = (xn.float() - xn_ref).abs()
    assert (e_xn > 1e-2 * xn_ref.abs() + 1e-2).sum().item() == 0, e_xn.max().item()
    e_t = (t.float() - t_ref).abs()
    bad = (e_t > 1e-2 * t_ref.abs() + 4e-2).sum().item()
    assert bad == 0, f'{bad} mismatches, max err {e_t.max().item():.4g}'
    assert e_t.mean().item() < 8e-3, e_t.mean().item()
