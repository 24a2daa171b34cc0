"""Stage-1 training path (first half): forward value and parameter gradients of
superresolution_b200.training.rrdbnet_forward (libsrb kernels for forward, data / weight / bias gradients and the
element-wise backward) against PyTorch autograd over the fp32 oracle RRDBNet with the same weights and input.
Forward: PSNR >= 50 dB against the plain fp32 oracle.  Gradients: against autograd over the oracle's arithmetic with the
kernels' rounding points (bf16 operands / stored activations, fp32 trunk) — at random init 0.05 % of the LeakyReLU
inputs change sign between a bf16 and an fp32 forward, which alone moves every gradient tensor by ~2.5 % per activation
layer; with matching signs the libsrb gradients (bf16-stored, fp32 trunk gradient) must agree to <= 4 % per tensor,
median <= 1.5 %."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


from _emulate import _emulated_rrdbnet  # noqa: E402  (oracle arithmetic with the kernels' rounding points)


def test_rrdbnet_forward_backward_matches_oracle_autograd(cuda):
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    from superresolution_b200.training import rrdbnet_forward
    torch.manual_seed(0)
    o = Oracle(smoothing='none').eval()
    m = Model(smoothing='none')
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda)
    g = torch.Generator().manual_seed(5)
    x = torch.randint(0, 65536, (2, 1, 32, 64), generator=g).float() / 65535.0
    R = torch.randn(2, 1, 64, 128, generator=g)
    # oracle autograd
    for p in o.stage1.parameters():
        p.grad = None
    with torch.no_grad():
        y_fp32 = o.stage1(x)          # the plain fp32 oracle: forward bar (PSNR >= 50 dB)
    yo = _emulated_rrdbnet(o.stage1, x)  # same arithmetic with the kernels' rounding points: gradient reference
    assert (yo.detach() - y_fp32).abs().max().item() < 2e-2
    (yo * R).sum().backward()
    # libsrb
    y = rrdbnet_forward(m.stage1, x.to(cuda))
    assert y.shape == yo.shape and y.requires_grad
    mse = (y.detach().cpu().double() - y_fp32.double()).pow(2).mean().item()
    psnr = 10 * math.log10(1.0 / max(mse, 1e-20))
    print(f'forward psnr {psnr:.2f} dB')
    assert psnr >= 50.0
    (y * R.to(cuda)).sum().backward()
    ref = dict(o.stage1.named_parameters())
    errs = {}
    for name, p in m.stage1.named_parameters():
        assert p.grad is not None, name
        assert p.grad.shape == ref[name].grad.shape
        errs[name] = _rel(p.grad.cpu(), ref[name].grad)
    for name in ('conv_last.weight', 'conv_last.bias', 'conv_hr.weight', 'conv_up1.bias', 'conv_body.weight',
                 'body.22.rdb3.conv5.weight', 'body.11.rdb2.conv3.weight', 'body.0.rdb1.conv1.weight',
                 'body.0.rdb1.conv1.bias', 'conv_first.weight', 'conv_first.bias'):
        print(f'{name:32s} rel L2 err {errs[name]:.4f}  |g| {ref[name].grad.norm().item():.4g}')
    v = sorted(errs.values())
    print(f'gradient tensors {len(v)}: median rel err {v[len(v) // 2]:.4f}, max {v[-1]:.4f} ({max(errs, key=errs.get)})')
    assert v[-1] <= 4e-2 and v[len(v) // 2] <= 1.5e-2


def test_rrdbnet_training_step_reduces_loss(cuda):
    """A few fused-optimizer steps on one batch through the libsrb forward/backward lower the reference loss."""
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    from superresolution_b200.losses import CombinedLoss
    from superresolution_b200.optim import FusedAdamW
    from superresolution_b200.training import rrdbnet_forward
    torch.manual_seed(1)
    m = Model(smoothing='none').to(cuda)
    opt = FusedAdamW(m.stage1.parameters(), lr=2e-4, weight_decay=0.0)
    crit = CombinedLoss()
    g = torch.Generator().manual_seed(6)
    x = (torch.randint(0, 65536, (2, 1, 32, 64), generator=g).float() / 65535.0).to(cuda)
    t = torch.nn.functional.interpolate(x, scale_factor=2, mode='bilinear', align_corners=False)
    losses = []
    for _ in range(6):
        opt.zero_grad()
        loss, _ = crit(rrdbnet_forward(m.stage1, x, accumulate_into_grad=True), t)
        loss.backward()
        opt.step(grad_scale=1.0, max_norm=1.0)
        losses.append(loss.item())
    print('losses', [round(v, 4) for v in losses])
    assert losses[-1] < losses[0]
