"""Parity at the production geometry (BASELINE config 2: 128x128 patches, batch 16).  The small-shape tests in
test_model_gpu.py take the per-layer conv path in stage 1; these shapes run what bench.py runs: the RRDB chain
kernel (trunk 64x64 tiles as 16x32), 2048 HAB tiles per launch, 64 images in the MLP's per-image scale table.
  * one full-size patch against the fp32 CPU oracle (PSNR >= 50 dB, the north-star bar for the bf16 path);
  * stage 1 alone on 64x64 inputs (smallest trunk the chain kernel accepts) against the oracle;
  * batch invariance at full size: every patch of a batch-16 forward equals the same patch run alone, bit for bit
    (a size-independent property: tiles, windows and the CAB pool never mix images, reductions are fixed-order)."""
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
    return o, m.to(cuda).eval()


def _input(B, H, W, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 65536, (B, 1, H, W), generator=g).float() / 65535.0


def test_stage1_chain_path_vs_oracle(pair, cuda):
    o, m = pair
    x = _input(2, 64, 64, seed=11)
    with torch.no_grad():
        ref = o.stage1(x)
        got = m.run(x.to(cuda), stage_mask=1).cpu()
    assert got.shape == ref.shape == (2, 1, 128, 128)
    p = psnr(got, ref)
    print(f'stage1 (chain kernel) psnr {p:.2f} dB')
    assert p >= 50.0


def test_full_size_patch_vs_oracle(pair, cuda):
    o, m = pair
    x = _input(1, 128, 128, seed=12)
    with torch.no_grad():
        ref = o(x)
        got = m(x.to(cuda)).cpu()
    assert got.shape == ref.shape == (1, 1, 512, 512)
    p = psnr(got, ref)
    print(f'full-size hybrid psnr {p:.2f} dB max-abs {(got - ref).abs().max().item():.4g}')
    assert p >= 50.0


def test_batch_invariance_full_size(pair, cuda):
    _, m = pair
    x = _input(16, 128, 128, seed=13).to(cuda)
    with torch.no_grad():
        full = m(x)
        assert full.shape == (16, 1, 512, 512)
        assert torch.isfinite(full).all()
        for i in (0, 7, 15):
            one = m(x[i:i + 1].contiguous())
            assert torch.equal(one[0], full[i]), f'patch {i}: max diff {(one[0] - full[i]).abs().max().item()}'
        again = m(x)
    assert torch.equal(again, full)  # run-to-run determinism (no atomics, fixed-order reductions)
