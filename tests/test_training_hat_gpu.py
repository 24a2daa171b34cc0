"""Stage-2 training path (second half): forward value, parameter gradients and input gradient of
superresolution_b200.training_hat.hat_forward (libsrb kernels for every forward op, data / weight / bias gradient,
LayerNorm / GELU / attention / channel-attention backward) against PyTorch autograd over the fp32 oracle HAT
(oracle/hat.py) with the same weights and input.  bf16 activations and activation gradients on the libsrb side: the
bars are relative L2 errors per tensor."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize('drop', [False, True])
def test_hat_forward_backward_matches_oracle_autograd(cuda, drop, monkeypatch):
    import oracle.hat as ohat
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    from superresolution_b200.training_hat import drop_path_masks, hat_forward
    torch.manual_seed(1)
    o = Oracle(smoothing='none')
    m = Model(smoothing='none')
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda)
    hat = o.stage2
    g = torch.Generator().manual_seed(7)
    B = 2
    x = torch.rand(B, 1, 32, 64, generator=g)
    R = torch.randn(B, 1, 64, 128, generator=g)
    masks = [(None, None)] * 24
    if drop:
        hat.train()
        torch.manual_seed(3)
        masks = [(None, None)] + [(drop_path_masks(B, 0.4, True, 'cpu'), drop_path_masks(B, 0.4, True, 'cpu'))
                                  for _ in range(23)]
        assert any(v == 0.0 for mm in masks[1:] for half in mm for v in half)
        it = iter([half for mm in masks[1:] for half in mm])

        def fake_drop_path(t, p=0.0, training=False):
            f = torch.as_tensor(next(it), dtype=t.dtype).view(-1, *([1] * (t.ndim - 1)))
            return t * f

        monkeypatch.setattr(ohat, 'drop_path', fake_drop_path)
    else:
        hat.eval()
    # libsrb
    xs = x.to(cuda).requires_grad_(True)
    dbg = {}
    y = hat_forward(m.stage2, xs, training=drop, masks=masks, debug=dbg)
    # oracle autograd.  The one LeakyReLU(0.01) of the network takes its sign pattern from the libsrb forward: ~0.3 % of
    # its inputs change sign between a bf16 and an fp32 forward, and a flipped element changes its gradient 100x, which
    # alone moves every upstream gradient tensor by ~8 % (measured) and says nothing about the backward kernels.
    pos = (dbg['u'].float() > 0).permute(0, 3, 1, 2).cpu()

    class SignMatchedLeakyReLU(torch.nn.Module):
        def forward(self, t):
            return t * torch.where(pos, 1.0, 0.01)

    hat.conv_before_upsample[1] = SignMatchedLeakyReLU()
    for p in hat.parameters():
        p.grad = None
    xo = x.clone().requires_grad_(True)
    yo = hat(xo)
    (yo * R).sum().backward()
    assert y.shape == yo.shape and y.requires_grad
    fe = _rel(y.detach().cpu(), yo.detach())
    print(f'forward rel err {fe:.4f}')
    assert fe <= 2e-2
    (y * R.to(cuda)).sum().backward()
    ref = dict(hat.named_parameters())
    errs = {}
    for n, p in m.stage2.named_parameters():
        assert p.grad is not None, n
        assert p.grad.shape == ref[n].grad.shape, n
        errs[n] = _rel(p.grad.cpu(), ref[n].grad)
    v = sorted(errs.values())
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:8]
    print('grad rel err: median %.4f  p90 %.4f  max %.4f' % (v[len(v) // 2], v[int(len(v) * 0.9)], v[-1]))
    print('worst:', worst)
    print('by depth:', [(n, round(e, 4)) for n, e in errs.items() if n.endswith('.weight') and ('blocks' not in n or 'proj' in n)])
    ex = _rel(xs.grad.cpu(), xo.grad)
    print(f'input grad rel err {ex:.4f}')
    assert v[len(v) // 2] <= 0.03
    assert v[-1] <= 0.15  # the worst tensors are the near-zero gradients of late bias terms; 0.118-0.121 measured
    assert ex <= 0.05
