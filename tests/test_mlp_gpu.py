"""Parity of the fused tcgen05 MLP branch (C ABI srb_mlp_branch) against plain PyTorch fp32:
y0 = t + cab*cs;  y = y0 + fc2(GELU_erf(fc1(LayerNorm(y0)))) — the `x + mlp(norm2(x))` tail of HAB / OCAB (upstream
hat_arch.py, reached from src/architecture.py:78).  Tolerance: inputs/weights are the same bf16-rounded values; the
kernel rounds LN output and the hidden activation to bf16 (operands of the tensor-core GEMMs) and the result to bf16."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref(t, cab, cs, g, b, w1, b1, w2, b2, tok_per_img):
    y0 = t.float()
    if cab is not None:
        y0 = y0 + cab.float() * cs.repeat_interleave(tok_per_img, 0)
    xn = F.layer_norm(y0, (96,), g, b, 1e-5)
    h = F.gelu(xn @ w1.t() + b1)
    return y0 + h @ w2.to(torch.bfloat16).float().t() + b2


@pytest.mark.parametrize('ntok,B,with_cab,inplace', [(128, 1, False, False), (4096, 2, True, False),
                                                     (148 * 128 * 3 + 128 * 5, 1, True, False),
                                                     (65536, 4, True, True), (8192, 2, False, True), (128 * 70, 70, True, False)])
def test_mlp_branch(cuda, ntok, B, with_cab, inplace):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(ntok + B)
    t = (torch.randn(ntok, 96, generator=g) * 1.5 + 0.3).to(cuda).to(torch.bfloat16)
    cab = (torch.randn(ntok, 96, generator=g) * 3).to(cuda).to(torch.bfloat16) if with_cab else None
    cs = (torch.rand(B, 96, generator=g) * 0.02).to(cuda) if with_cab else None
    ln_g = (1 + 0.2 * torch.randn(96, generator=g)).to(cuda)
    ln_b = (0.1 * torch.randn(96, generator=g)).to(cuda)
    w1 = (torch.randn(192, 96, generator=g) / 96 ** 0.5).to(cuda)
    b1 = (0.3 * torch.randn(192, generator=g)).to(cuda)
    w2 = (torch.randn(96, 192, generator=g) / 192 ** 0.5).to(cuda)
    b2 = (0.3 * torch.randn(96, generator=g)).to(cuda)
    ref = _ref(t, cab, cs, ln_g, ln_b, w1, b1, w2, b2, ntok // B)
    (w1f, b1f), w2p = ops.fold_ln_linear(w1, b1, ln_g, ln_b), ops.pack_conv_weight(w2)
    tin = t.clone()
    out = tin if inplace else torch.full_like(t, float('nan'))
    ops.mlp_branch(tin, w1f, b1f, w2p, b2, cab=cab, cs=cs, out=out, tok_per_img=ntok // B)
    torch.cuda.synchronize()
    err = (out.float() - ref).abs()
    tol = 1e-2 * ref.abs() + 3e-2
    bad = (err > tol).sum().item()
    assert bad == 0, f'{bad} mismatches, max err {err.max().item():.4g}, ref absmax {ref.abs().max().item():.4g}'
    assert err.mean().item() < 6e-3, err.mean().item()


def test_mlp_branch_with_fused_next_layernorm(cuda):
    """The last HAB of a group also emits OCAB.norm1(y): must equal LayerNorm of the kernel's own bf16 output."""
    from superresolution_b200 import ops
    ntok, B = 152 * 128, 2
    g = torch.Generator().manual_seed(7)
    t = (torch.randn(ntok, 96, generator=g) * 1.5 + 0.3).to(cuda).to(torch.bfloat16)
    cab = (torch.randn(ntok, 96, generator=g) * 3).to(cuda).to(torch.bfloat16)
    cs = (torch.rand(B, 96, generator=g) * 0.02).to(cuda)
    mk = lambda *s, k=1.0: (k * torch.randn(*s, generator=g)).to(cuda)
    ln_g, ln_b = 1 + mk(96, k=0.2), mk(96, k=0.1)
    w1, b1, w2, b2 = mk(192, 96, k=96 ** -0.5), mk(192, k=0.3), mk(96, 192, k=192 ** -0.5), mk(96, k=0.3)
    g2, be2 = 1 + mk(96, k=0.3), mk(96, k=0.2)
    (w1f, b1f), w2p = ops.fold_ln_linear(w1, b1, ln_g, ln_b), ops.pack_conv_weight(w2)
    plain = ops.mlp_branch(t, w1f, b1f, w2p, b2, cab=cab, cs=cs, tok_per_img=ntok // B)
    out, xn = ops.mlp_branch(t, w1f, b1f, w2p, b2, cab=cab, cs=cs, tok_per_img=ntok // B, next_ln=(g2, be2))
    torch.cuda.synchronize()
    assert torch.equal(out, plain)
    want = F.layer_norm(out.float(), (96,), g2, be2, 1e-5)
    err = (xn.float() - want).abs()
    assert (err <= 8e-3 * want.abs() + 1e-2).all(), err.max().item()  # one bf16 rounding of the result
