"""Backward of the conv / linear layers (SURVEY.md §8 a12 building blocks) against torch autograd of F.conv2d on the
same bf16-rounded operands (fp32 math in the reference).  Data gradient = the forward tcgen05 kernel on flipped
weights (C ABI srb_pack_conv_weight_dgrad + srb_conv_nhwc_bf16), weight / bias gradient = srb_conv_wgrad."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# B, H, W, Cin, Cout, taps
CASES = [(1, 32, 32, 64, 32, 9), (2, 32, 64, 96, 32, 9), (1, 64, 64, 160, 32, 9), (2, 32, 32, 192, 64, 9),
         (1, 32, 32, 32, 96, 9), (2, 64, 32, 96, 96, 9), (1, 64, 64, 64, 256, 9), (2, 32, 32, 96, 288, 1),
         (1, 64, 64, 192, 96, 1), (3, 32, 32, 64, 64, 9), (1, 32, 64, 96, 192, 1), (2, 64, 64, 96, 96, 1),
         (1, 16, 32, 96, 288, 1), (1, 32, 32, 64, 96, 1)]
# the tcgen05 weight-gradient kernel takes ragged chunks (TMA zero fill) and a full 128-channel M chunk; the last two
# shapes (Cin 16, Cout 8: conv_first / the padded conv_last) stay on the mma.sync kernels
WGRAD_EXTRA = [(1, 6, 40, 64, 32, 9), (2, 10, 24, 32, 96, 9), (1, 5, 7, 96, 96, 1), (1, 32, 32, 128, 32, 9),
               (1, 32, 32, 16, 96, 9), (1, 32, 32, 64, 8, 9)]


def _mk(B, H, W, Cin, Cout, taps, cuda, seed):
    g = torch.Generator().manual_seed(seed)
    k = 3 if taps == 9 else 1
    x = torch.randn(B, H, W, Cin, generator=g).to(cuda).bfloat16()
    dy = torch.randn(B, H, W, Cout, generator=g).to(cuda).bfloat16()
    w = (torch.randn(Cout, Cin, k, k, generator=g) * (Cin * taps) ** -0.5).to(cuda)
    return x, dy, w, k


@pytest.mark.parametrize('B,H,W,Cin,Cout,taps', CASES)
def test_conv_dgrad(cuda, B, H, W, Cin, Cout, taps):
    from superresolution_b200 import ops
    x, dy, w, k = _mk(B, H, W, Cin, Cout, taps, cuda, 1)
    wq = w.bfloat16().float()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    F.conv2d(xr, wq, padding=k // 2).backward(dy.float().permute(0, 3, 1, 2))
    want = xr.grad.permute(0, 2, 3, 1)
    got = torch.empty(B, H, W, Cin, device=cuda, dtype=torch.bfloat16)
    ops.conv_nhwc(dy, Cout, ops.pack_conv_weight_dgrad(w), None, out=got)
    err = (got.float() - want).abs()
    assert (err <= 1e-2 * want.abs() + 2e-2).all(), err.max().item()
    # accumulation into an existing gradient slice (dense-block concat): G += dgrad, in place
    G = torch.randn(B, H, W, Cin + 32, device=cuda).bfloat16()
    G0 = G.clone()
    ops.conv_nhwc(dy, Cout, ops.pack_conv_weight_dgrad(w), None, out=G[..., :Cin], res1=G[..., :Cin])
    err = (G[..., :Cin].float() - (G0[..., :Cin].float() + want)).abs()
    assert (err <= 1e-2 * want.abs() + 4e-2).all(), err.max().item()
    assert torch.equal(G[..., Cin:], G0[..., Cin:])


@pytest.mark.parametrize('B,H,W,Cin,Cout,taps', CASES + WGRAD_EXTRA)
def test_conv_wgrad_and_bias_grad(cuda, B, H, W, Cin, Cout, taps):
    from superresolution_b200 import ops
    x, dy, w, k = _mk(B, H, W, Cin, Cout, taps, cuda, 2)
    wr = w.clone().requires_grad_(True)
    br = torch.zeros(Cout, device=cuda, requires_grad=True)
    F.conv2d(x.float().permute(0, 3, 1, 2), wr, br, padding=k // 2).backward(dy.float().permute(0, 3, 1, 2))
    # operands as channel-offset views of wider buffers (dense-block layout)
    xb = torch.zeros(B, H, W, Cin + 32, device=cuda, dtype=torch.bfloat16)
    xb[..., :Cin] = x
    dyb = torch.zeros(B, H, W, Cout + 64, device=cuda, dtype=torch.bfloat16)
    dyb[..., 32:32 + Cout] = dy
    dw, db = ops.conv_wgrad(xb[..., :Cin], Cin, dyb[..., 32:32 + Cout], Cout, taps)
    scale = (B * H * W) ** 0.5  # magnitude of a sum of B*H*W unit-variance products
    assert (dw.view_as(wr.grad) - wr.grad).abs().max().item() <= 2e-3 * scale
    assert (db - br.grad).abs().max().item() <= 2e-3 * scale
    dw2, db2 = ops.conv_wgrad(xb[..., :Cin], Cin, dyb[..., 32:32 + Cout], Cout, taps, dw=dw.clone(), db=db.clone(),
                              accumulate=True)
    assert torch.allclose(dw2, 2 * dw, rtol=1e-6, atol=1e-6) and torch.allclose(db2, 2 * db, rtol=1e-6, atol=1e-6)
    dw3, _ = ops.conv_wgrad(xb[..., :Cin], Cin, dyb[..., 32:32 + Cout], Cout, taps)
    assert torch.equal(dw3, dw)  # deterministic
