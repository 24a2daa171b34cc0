"""HAT (stage 2) backward building blocks against torch autograd in fp32 on the same bf16-rounded inputs:
LayerNorm(96) forward / backward (+ fused shortcut gradient), exact-erf GELU forward / backward.
(tested operators for the stage-2 training path, which is not assembled yet.)"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('ntok,with_res', [(64, None), (4096 + 37, 'bf16'), (148 * 4 * 16 * 3 + 5, 'f32')])
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
    assert (y.float() - y_ref.detach()).abs().max().item() <= 3e-2
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
