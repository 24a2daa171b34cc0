"""Parity of the libsrb forward (through the drop-in nn.Module) against the CPU fp32 oracle on the same seeded
weights and inputs.  Bar (BASELINE.json north_star): bf16 path PSNR(output, fp32 oracle) >= 50 dB on the [0,1] scale.
Stage-level checks localise a failure: RRDBNet only, HAT only, full hybrid (with and without smoothing)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def psnr(a, b):
    mse = (a.double() - b.double()).pow(2).mean().item()
    return 10 * math.log10(1.0 / max(mse, 1e-20))


@pytest.fixture(scope='module')
def pair(cuda):
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    torch.manual_seed(0)
    o = Oracle(smoothing='balanced').eval()
    m = Model(smoothing='balanced')
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda).eval()
    return o, m


def _input(B, H, W, seed=1):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 65536, (B, 1, H, W), generator=g).float() / 65535.0


def test_stage1_rrdbnet(pair, cuda):
    o, m = pair
    x = _input(2, 32, 32)
    with torch.no_grad():
        ref = o.stage1(x)
        got = m.run(x.to(cuda), stage_mask=1).cpu()
    assert got.shape == ref.shape == (2, 1, 64, 64)
    p = psnr(got, ref)
    print(f'stage1 psnr {p:.2f} dB max-abs {(got - ref).abs().max().item():.4g}')
    assert p >= 50.0


def test_stage2_hat(pair, cuda):
    o, m = pair
    x = _input(2, 64, 64, seed=2)
    with torch.no_grad():
        ref = o.stage2(x)
        got = m.run(x.to(cuda), stage_mask=2).cpu()
    assert got.shape == ref.shape == (2, 1, 128, 128)
    p = psnr(got, ref)
    print(f'stage2 psnr {p:.2f} dB max-abs {(got - ref).abs().max().item():.4g}')
    assert p >= 50.0


def test_full_hybrid_balanced(pair, cuda):
    o, m = pair
    x = _input(1, 32, 32, seed=3)
    with torch.no_grad():
        o.output_size = 128
        ref = o(x)
        m.output_size = 128
        got = m(x.to(cuda)).cpu()
    assert got.shape == ref.shape == (1, 1, 128, 128)
    p = psnr(got, ref)
    print(f'full psnr {p:.2f} dB max-abs {(got - ref).abs().max().item():.4g}')
    assert p >= 50.0


def test_u16_input_and_output(pair, cuda):
    """u16 input is divided by 65535 (src/dataset.py:34); u16 output is clip*65535 truncated (Modello_5.py:29-33)."""
    o, m = pair
    g = torch.Generator().manual_seed(4)
    xi = torch.randint(0, 65536, (1, 1, 32, 32), generator=g)
    xf = xi.float() / 65535.0
    m.output_size = 128
    with torch.no_grad():
        a = m(xf.to(cuda))
        b = m.run(xi.to(torch.uint16).to(cuda), 3)
        c = m.run(xf.to(cuda), 3, out_dtype=torch.uint16)
    assert torch.equal(a, b)
    want = (a.clamp(0, 1) * 65535).to(torch.int32)
    assert torch.equal(c.to(torch.int32), want)
