"""Every geometry the reference accepts: H, W multiples of 4 — pixel_unshuffle(2) in stage 1, the
8 x 8 HAT windows at 2H x 2W in stage 2 (src/architecture.py:74-82) — including sizes that are ragged against every
kernel tile (conv 16 x 16/32, conv_cout1 8 x 32, MLP 128 tokens, HAB window pairs).  Same bars as the whole-tile
geometries: PSNR >= 50 dB against the fp32 oracle, and with the trained-like fixture rel-L2 <= REL_BAR per stage."""
import math

import pytest
import torch

from _trained_like import make_trained_like

pytestmark = pytest.mark.gpu


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


@pytest.mark.parametrize('B,H,W', [(1, 36, 44), (2, 100, 100), (1, 4, 4), (3, 20, 68), (1, 52, 32)])
def test_stage1_ragged(pair, cuda, B, H, W):
    o, m = pair
    x = _input(B, H, W, seed=H * 1000 + W)
    with torch.no_grad():
        ref = o.stage1(x)
        got = m.run(x.to(cuda), stage_mask=1).cpu()
    assert got.shape == ref.shape == (B, 1, 2 * H, 2 * W)
    r = rel_l2(got, ref)
    print(f'stage1 {B}x{H}x{W}: rel-L2 {r:.4f}')
    assert torch.isfinite(got).all() and r <= 0.02


@pytest.mark.parametrize('B,H,W', [(1, 72, 88), (2, 8, 8), (1, 40, 24), (3, 24, 56)])
def test_stage2_ragged(pair, cuda, B, H, W):
    o, m = pair
    x = _input(B, H, W, seed=H * 1000 + W + 1)
    with torch.no_grad():
        ref = o.stage2(x)
        got = m.run(x.to(cuda), stage_mask=2).cpu()
    assert got.shape == ref.shape == (B, 1, 2 * H, 2 * W)
    r = rel_l2(got, ref)
    print(f'stage2 {B}x{H}x{W}: rel-L2 {r:.4f}')
    assert torch.isfinite(got).all() and r <= 0.02


@pytest.mark.parametrize('B,H,W', [(1, 36, 44), (1, 100, 100), (2, 12, 20)])
def test_full_hybrid_ragged(pair, cuda, B, H, W):
    """Non-square frames; the B = 2 case asks for a 64 x 64 output, so the reference's bicubic branch runs too
    (src/architecture.py:80-81: taken when the last dimension differs from output_size)."""
    o, m = pair
    x = _input(B, H, W, seed=H * 1000 + W + 2)
    o.output_size = m.output_size = 64 if B == 2 else 4 * W
    with torch.no_grad():
        ref = o(x)
        got = m(x.to(cuda)).cpu()
    assert got.shape == ref.shape
    p, r = psnr(got, ref), rel_l2(got, ref)
    print(f'hybrid {B}x{H}x{W} -> {tuple(got.shape)}: psnr {p:.2f} dB  rel-L2 {r:.4f}')
    assert torch.isfinite(got).all() and p >= 50.0 and r <= 0.05


def test_random_init_ragged_frame_psnr(cuda):
    """the north-star bar on the default initialisation at 100 x 100 (BASELINE: PSNR >= 50 dB vs the fp32 reference)"""
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    torch.manual_seed(3)
    o = Oracle(output_size=400, smoothing='balanced').eval()
    m = Model(output_size=400, smoothing='balanced')
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda).eval()
    x = _input(1, 100, 100, seed=9)
    with torch.no_grad():
        ref = o(x)
        got = m(x.to(cuda)).cpu()
    assert got.shape == ref.shape == (1, 1, 400, 400)
    assert psnr(got, ref) >= 50.0


@pytest.mark.parametrize('B,H,W', [(1, 36, 44), (2, 20, 68), (1, 100, 100), (2, 64, 64)])
def test_forward_stays_inside_its_buffers(pair, cuda, B, H, W):
    """Guard bands around the workspace, the input and the output of srb_forward (C ABI, caller-owned buffers): the
    ragged-tile paths (TMA-clipped stores, masked epilogues, clamped MLP rows) and the whole-tile ones must not write a
    byte outside what srb_workspace_bytes / the output shape promise, and the result equals the module's."""
    from superresolution_b200 import _lib as L
    _, m = pair
    lib = L.lib()
    G = 1 << 16  # guard bytes on either side
    x = _input(B, H, W, seed=7).to(cuda)
    m.output_size = 4 * W
    with torch.no_grad():
        want = m.run(x, stage_mask=3)
    packed = m.pack_weights()
    nws = lib.srb_workspace_bytes(B, H, W)
    nout = B * 16 * H * W * 4
    arena = torch.full((G + nws + G + nout + G,), 0xA5, dtype=torch.uint8, device=cuda)
    ws = arena[G:G + nws]
    y = arena[G + nws + G:G + nws + G + nout]
    assert ws.data_ptr() % 256 == 0 and y.data_ptr() % 16 == 0
    L.check(lib.srb_forward(L.ptr(x), L.SRB_F32, L.ptr(y), L.SRB_F32, B, H, W, L.ptr(packed), L.ptr(ws), nws, m._smoothing,
                            3, L.current_stream()), 'forward')
    torch.cuda.synchronize()
    for lo, hi in ((0, G), (G + nws, G + nws + G), (G + nws + G + nout, G + nws + G + nout + G)):
        assert bool((arena[lo:hi] == 0xA5).all()), (lo, hi)
    got = y.view(torch.float32).view(B, 1, 4 * H, 4 * W)
    assert torch.equal(got, want)
