"""End-to-end parity with 'trained-like' parameters (tests/_trained_like.py).

At random init the final output has sigma ~0.015, so the north-star bar (PSNR >= 50 dB on the [0,1] scale) would accept
a 20 % relative error; with every bias, LayerNorm affine, relative-position table, channel attention and the CAB
branch switched on, each of those paths moves the output by >= 4 % (measured with the oracle: zeroing the SA bias
tables 16 %, the OCAB tables 4 %, cab.0 5.5 %, channel attention 4.4 %, one RRDB 16 %), and the assertion is a
signal-relative one:  ||got - ref||_2 / ||ref - mean(ref)||_2  <=  REL_BAR  in addition to the 50 dB bar.
Reference path: src/architecture.py:74-82 (stage1 -> s1 -> stage2 -> s2 -> sf) through oracle/ (fp32, CPU)."""
import math

import pytest
import torch

from _trained_like import make_trained_like

pytestmark = pytest.mark.gpu

# bf16 path, relative L2 against the fp32 oracle.  Measured on B200: each stage alone 1.0 - 1.1 %; the full
# hybrid 3.2 - 3.5 % (the stage-1 error is the input of a 24-block transformer and passes two blurs): bars ~1.5 - 2x that.
REL_BAR = 0.02
REL_BAR_HYBRID = 0.05


def psnr(a, b):
    mse = (a.double() - b.double()).pow(2).mean().item()
    return 10 * math.log10(1.0 / max(mse, 1e-20))


def rel_l2(got, ref):
    ref = ref.double()
    return ((got.double() - ref).norm() / (ref - ref.mean()).norm()).item()


@pytest.fixture(scope='module')
def pair(cuda):
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    torch.manual_seed(0)
    o = make_trained_like(Oracle(smoothing='balanced').eval())
    m = Model(smoothing='balanced')
    m.load_state_dict(o.state_dict(), strict=True)
    return o, m.to(cuda).eval()


def _input(B, H, W, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 65536, (B, 1, H, W), generator=g).float() / 65535.0


def _check(tag, got, ref, bar=REL_BAR, psnr_bar=None):
    p, r = psnr(got, ref), rel_l2(got, ref)
    print(f'{tag}: sigma(ref) {ref.std().item():.4f}  psnr {p:.2f} dB  rel-L2 {r:.4f}  max-abs {(got - ref).abs().max().item():.4g}')
    assert torch.isfinite(got).all()
    assert r <= bar, r
    if psnr_bar is not None:  # the north-star bar is defined on the [0,1] model output: asserted on the full hybrid only
        assert p >= psnr_bar, p   # (the amplified stage outputs leave [0,1]: sigma 0.25 - 0.6)


@pytest.mark.parametrize('hw', [32, 64])
def test_stage1_trained_like(pair, cuda, hw):
    o, m = pair
    x = _input(2, hw, hw, seed=21 + hw)
    with torch.no_grad():
        ref = o.stage1(x)
        got = m.run(x.to(cuda), stage_mask=1).cpu()
    _check(f'stage1 {hw}', got, ref)


@pytest.mark.parametrize('hw', [64, 128])
def test_stage2_trained_like(pair, cuda, hw):
    o, m = pair
    x = _input(1, hw, hw, seed=31 + hw)
    with torch.no_grad():
        ref = o.stage2(x)
        got = m.run(x.to(cuda), stage_mask=2).cpu()
    _check(f'stage2 {hw}', got, ref)


@pytest.mark.parametrize('hw', [32, 64, 128])
def test_full_hybrid_trained_like(pair, cuda, hw):
    o, m = pair
    x = _input(1, hw, hw, seed=41 + hw)
    o.output_size = m.output_size = 4 * hw
    with torch.no_grad():
        ref = o(x)
        got = m(x.to(cuda)).cpu()
    assert got.shape == ref.shape == (1, 1, 4 * hw, 4 * hw)
    assert ref.std().item() >= 0.05  # the fixture keeps the final signal well above the random-init 0.015
    _check(f'hybrid {hw}', got, ref, bar=REL_BAR_HYBRID, psnr_bar=50.0)


def test_branches_are_live(pair):
    """Guards the fixture itself (CPU oracle only): killing a branch must move the output of the stage it belongs to by
    at least twice the stage bar, so that the stage tests above would catch it."""
    import copy
    o, _ = pair
    x1, x2 = _input(1, 32, 32, seed=51), _input(1, 64, 64, seed=52)
    with torch.no_grad():
        ref1, ref2 = o.stage1(x1), o.stage2(x2)
        for pat in ('attn.relative_position_bias_table', 'overlap_attn.relative_position_bias_table', 'cab.0.weight',
                    'attn.qkv.bias', 'stage1.body.22.'):
            o2 = copy.deepcopy(o)
            for n, p in o2.named_parameters():
                if pat in n:
                    p.zero_()
            d = rel_l2(o2.stage1(x1), ref1) if pat.startswith('stage1') else rel_l2(o2.stage2(x2), ref2)
            print(f'zero {pat}: stage output moves by {d:.3f}')
            assert d >= 2 * REL_BAR, (pat, d)
