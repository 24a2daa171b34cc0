"""GPU parity of the HBM-bound neighbours (C ABI: srb_weighted_l1, srb_psnr_ssim, srb_log_normalize_u16,
srb_blend_*) against the oracle restatements of the reference's own code and the committed reference goldens.
Tolerances: loss rel 1e-5; PSNR/SSIM 1e-4 (north_star); LOG-normalise bit-exact except log1pf/np.log1p 1-ulp
differences that flip a truncation (<= 1 LSB, <= 0.1% of pixels); blend fp32 1e-5."""
import os

import numpy as np
import pytest
import torch

from _inputs import aux_inputs

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_weighted_l1_value_and_grad(cuda):
    from oracle import aux
    from superresolution_b200.losses import CombinedLoss
    g = np.load(os.path.join(HERE, 'golden', 'aux.npz'))
    p, t, _ = aux_inputs()
    loss, d = CombinedLoss()(p.to(cuda), t.to(cuda))
    assert set(d) == {'total', 'weighted_l1'} and d['total'] is loss
    assert abs(loss.item() - float(g['loss'])) <= 1e-5 * abs(float(g['loss']))
    # gradient vs autograd through the oracle (reference src/losses.py expression)
    pc = p.clone().requires_grad_(True)
    aux.combined_loss(pc, t).backward()
    pg = p.to(cuda).requires_grad_(True)
    CombinedLoss()(pg, t.to(cuda))[0].backward()
    assert torch.allclose(pg.grad.cpu(), pc.grad, rtol=1e-6, atol=1e-9)
    # large, both branches, odd size (tail path)
    gen = torch.Generator().manual_seed(5)
    t2 = torch.rand(3, 1, 257, 131, generator=gen) * 0.04
    p2 = t2 + 0.01 * torch.randn(3, 1, 257, 131, generator=gen)
    got = CombinedLoss()(p2.to(cuda), t2.to(cuda))[0].item()
    ref = aux.combined_loss(p2, t2).item()
    assert abs(got - ref) <= 1e-5 * abs(ref)


def test_psnr_ssim(cuda):
    from oracle import aux
    from superresolution_b200.metrics import Metrics
    g = np.load(os.path.join(HERE, 'golden', 'aux.npz'))
    p, t, _ = aux_inputs()
    m = Metrics()
    m.update(p.to(cuda), t.to(cuda))
    r = m.compute()
    assert abs(r['psnr'] - float(g['psnr'])) < 1e-4 and abs(r['ssim'] - float(g['ssim'])) < 1e-4
    # 512x512 pairs as in config 5, values outside [0,1] to exercise the clamp; accumulate over two updates
    gen = torch.Generator().manual_seed(6)
    t2 = torch.rand(3, 1, 512, 512, generator=gen)
    p2 = t2 + 0.01 * torch.randn(3, 1, 512, 512, generator=gen)
    mo, mg = aux.Metrics(), Metrics()
    for a, b in ((p2, t2), (p2[:1] * 1.2 - 0.1, t2[:1])):
        mo.update(a, b)
        mg.update(a.to(cuda), b.to(cuda))
    ro, rg = mo.compute(), mg.compute()
    assert abs(ro['psnr'] - rg['psnr']) < 1e-4 and abs(ro['ssim'] - rg['ssim']) < 1e-4
    # known answers
    mg = Metrics()
    mg.update(t2.to(cuda), t2.to(cuda))
    r = mg.compute()
    assert abs(r['ssim'] - 1.0) < 1e-5 and abs(r['psnr'] - 80.0) < 1e-3


def test_backup_loss_terms_value_and_grad(cuda):
    """Charbonnier / astro terms (Backup/src/losses.py:6-15,37-48) on srb_charbonnier_astro against the
    reference-minted golden values and autograd through the oracle; odd size exercises the tail path."""
    from oracle import aux
    from superresolution_b200.losses import AstroCombinedLoss, CharbonnierLoss
    g = np.load(os.path.join(HERE, 'golden', 'backup_loss.npz'))
    p, t, _ = aux_inputs()
    total, d = AstroCombinedLoss(1.0, 0.0, 0.05)(p.to(cuda), t.to(cuda))
    assert abs(d['char'].item() - float(g['char'])) <= 1e-5 * float(g['char'])
    assert abs(d['astro'].item() - float(g['astro'])) <= 1e-5 * float(g['astro'])
    assert abs(d['l1_raw'].item() - float(g['l1_raw'])) <= 1e-5 * float(g['l1_raw'])
    assert abs(total.item() - (float(g['char']) + 0.05 * float(g['astro']))) <= 1e-5 * total.item()
    assert abs(CharbonnierLoss()(p.to(cuda), t.to(cuda)).item() - float(g['char'])) <= 1e-5 * float(g['char'])
    gen = torch.Generator().manual_seed(9)
    t2 = torch.rand(3, 1, 129, 67, generator=gen)
    p2 = t2 + 0.02 * torch.randn(3, 1, 129, 67, generator=gen)
    pc = p2.clone().requires_grad_(True)
    c, _, a = aux.astro_loss_terms(pc, t2)
    (0.7 * c + 0.3 * a).backward()
    pg = p2.to(cuda).requires_grad_(True)
    tot, _ = AstroCombinedLoss(0.7, 0.0, 0.3)(pg, t2.to(cuda))
    tot.backward()
    assert abs(tot.item() - (0.7 * c + 0.3 * a).item()) <= 1e-5 * tot.item()
    assert torch.allclose(pg.grad.cpu(), pc.grad, rtol=2e-4, atol=1e-9)
    with pytest.raises(NotImplementedError):
        AstroCombinedLoss(1.0, 0.05, 0.05)


def test_ssim_torch_drop_in_and_range_check(cuda):
    """ssim_torch mirrors src/metrics.py:5-19 for inputs in [0,1]; out-of-range inputs are refused (the reference
    function does not clamp, the fused kernel does: ADVICE r01)."""
    from oracle import aux
    from superresolution_b200.metrics import ssim_torch
    gen = torch.Generator().manual_seed(8)
    t2 = torch.rand(2, 1, 96, 80, generator=gen)
    p2 = (t2 + 0.05 * torch.randn(2, 1, 96, 80, generator=gen)).clamp(0, 1)
    got = ssim_torch(p2.to(cuda), t2.to(cuda)).item()
    assert abs(got - aux.ssim_torch(p2, t2).item()) < 1e-4
    with pytest.raises(ValueError, match='outside'):
        ssim_torch((p2 * 1.5).to(cuda), t2.to(cuda))


def test_log_normalize(cuda):
    from oracle import aux
    from superresolution_b200.normalization import log_normalize_u16
    g = np.load(os.path.join(HERE, 'golden', 'aux.npz'))
    _, _, raw = aux_inputs()
    got = log_normalize_u16(torch.from_numpy(raw).to(cuda), g['gmin'], g['gmax']).cpu().numpy()
    ref = g['lognorm_u16']
    diff = np.abs(got.astype(np.int64) - ref.astype(np.int64))
    assert diff.max() <= 1 and (diff > 0).mean() <= 1e-3
    assert got[3, 5] == 0 and got[7, 9] == 65535 and got[8, 9] == 0  # nan, +inf, -inf
    # config-5 sized pair (512x512 + 128x128), heavy-tailed counts
    rng = np.random.default_rng(7)
    big = np.expm1(rng.normal(1.0, 1.5, size=(512 * 512 + 128 * 128,))).astype(np.float32)
    big[rng.random(big.shape) < 0.001] = np.nan
    big[rng.random(big.shape) < 0.01] *= -1
    ref = aux.log_normalize_u16(big, 0.05, 6.0)
    got = log_normalize_u16(torch.from_numpy(big).to(cuda), 0.05, 6.0).cpu().numpy()
    diff = np.abs(got.astype(np.int64) - ref.astype(np.int64))
    assert diff.max() <= 1 and (diff > 0).mean() <= 1e-3
    # The float32 log1p of the reference platform (numpy -> libm log1pf) is itself NOT correctly rounded (glibc 2.39: 13.9 %
    # of float32 inputs are 1 ulp off), so "the reference's bits" are libm-specific.  What is asserted exactly: every
    # output pixel equals the reference's float32 pipeline (:167-179) evaluated with SOME log1p within 1 ulp of the
    # correctly rounded one (CUDA log1pf and numpy's both are), i.e. lies between the pipeline on CR -/+ 1 ulp.
    clean = np.maximum(np.nan_to_num(big), 0)
    l_cr = np.log1p(clean.astype(np.float64)).astype(np.float32)

    def pipe(l):
        d = (l - np.float32(0.05)) / (np.float32(6.0) - np.float32(0.05) + 1e-8)
        return (np.clip(d, 0, 1) * 65535).astype(np.uint16)
    lo, hi = pipe(np.nextafter(l_cr, np.float32(-np.inf))), pipe(np.nextafter(l_cr, np.float32(np.inf)))
    assert ((got >= lo) & (got <= hi)).all()
    assert ((ref >= lo) & (ref <= hi)).all()  # numpy's own result obeys the same statement
    assert (got == pipe(l_cr)).mean() >= 0.999
    # monotone + idempotent range properties, empty input
    xs = torch.linspace(0, 400, 100001, device=cuda)
    v = log_normalize_u16(xs, 0.05, 6.0).cpu().numpy().astype(np.int64)
    assert (np.diff(v) >= 0).all() and v[0] == 0 and v[-1] <= 65535
    assert log_normalize_u16(torch.zeros(0, device=cuda), 0.0, 1.0).numel() == 0


def test_tiled_frame_matches_oracle_blend(cuda):
    """ragged frame (clamped last row/col), tile outputs from the CUDA model, stitch vs the oracle blend of the SAME
    tile outputs; 2-band sharding reproduces the 1-band frame."""
    from oracle import tiling as otil
    from superresolution_b200 import tiling
    from superresolution_b200.architecture import HybridSuperResolutionModel
    torch.manual_seed(0)
    m = HybridSuperResolutionModel(output_size=128, smoothing='balanced').to(cuda).eval()
    g = torch.Generator().manual_seed(2)
    frame = torch.randint(0, 65536, (80, 56), generator=g).to(torch.uint16).to(cuda)
    tile, stride, scale = 32, 24, 4
    out, (r0, r1) = tiling.infer_frame(m, frame, tile, stride, batch=4, out_dtype=torch.float32)
    assert (r0, r1) == (0, 320) and tuple(out.shape) == (320, 224)
    origins = tiling.plan(80, 56, tile, stride)
    with torch.no_grad():
        tiles = [m.run(frame[y:y + tile, x:x + tile][None, None].contiguous(), 3)[0, 0].cpu().numpy()
                 for (y, x) in origins]
    ref = otil.blend(tiles, [(scale * y, scale * x) for (y, x) in origins], 320, 224, scale * (tile - stride))
    assert np.abs(out.cpu().numpy() - ref).max() < 1e-5
    # overlapping tiles reach a pixel in the plan's raster order (one tile per blend launch, no atomics): bit-reproducible
    for batch in (4, 3):
        again, _ = tiling.infer_frame(m, frame, tile, stride, batch=batch, out_dtype=torch.float32)
        assert torch.equal(again, out)
    bands = [tiling.infer_frame(m, frame, tile, stride, batch=4, rank=r, world=2, out_dtype=torch.float32,
                                seams='recompute')[0]
             for r in range(2)]
    assert np.abs(torch.cat(bands).cpu().numpy() - ref).max() < 1e-5
    # seams='p2p': every tile computed once, the overhanging partial sums handed to the next rank (here through an
    # in-process stand-in for the isend / irecv pair; the real exchange is covered by the gloo test on CPU)
    strips = {}

    def handover(sums, plan_all, rank, group=None):
        _, own0, own1, end = plan_all[rank]
        if rank + 1 < len(plan_all) and end > own1:
            strips[rank] = sums[:, own1 - own0:].clone()
        if rank - 1 in strips:
            sums[:, :strips[rank - 1].shape[1]] += strips[rank - 1]
        return sums

    parts = [tiling.infer_frame(m, frame, tile, stride, batch=4, rank=r, world=2, out_dtype=torch.float32, seams='p2p',
                                _exchange=handover) for r in range(2)]
    assert [p[1] for p in parts] == [(0, 96), (96, 320)] and strips[0].shape[1] == 32
    assert np.abs(torch.cat([p[0] for p in parts]).cpu().numpy() - ref).max() < 1e-5
    u16, _ = tiling.infer_frame(m, frame, tile, stride, batch=16)
    want = (np.clip(out.cpu().numpy(), 0, 1) * 65535).astype(np.uint16)
    assert np.abs(u16.cpu().numpy().astype(np.int64) - want.astype(np.int64)).max() <= 1


def test_model_matches_committed_golden(cuda):
    import math
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from _inputs import model_inputs
    g = np.load(os.path.join(HERE, 'golden', 'model.npz'))
    torch.manual_seed(0)
    o = Oracle(output_size=128, smoothing='balanced')
    m = HybridSuperResolutionModel(output_size=128, smoothing='balanced')
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda).eval()
    x32, x64 = model_inputs()
    with torch.no_grad():
        outs = {'stage1': m.run(x32.to(cuda), 1), 'stage2': m.run(x64.to(cuda), 2), 'full': m(x32.to(cuda))}
        m2 = HybridSuperResolutionModel(output_size=128, smoothing='none')
        m2.load_state_dict({k: v for k, v in o.state_dict().items() if not k.startswith(('s1.', 's2.', 'sf.'))})
        outs['full_none'] = m2.to(cuda).eval()(x32.to(cuda))
    for k, v in outs.items():
        mse = float(((v.cpu().double().numpy() - g[k].astype(np.float64)) ** 2).mean())
        psnr = 10 * math.log10(1.0 / max(mse, 1e-20))
        print(k, f'{psnr:.2f} dB')
        assert psnr >= 50.0, (k, psnr)


def _sky(g, n, shape):
    """positive heavy-tailed 'astronomical counts' with NaN, inf, negatives and exact zeros mixed in"""
    x = torch.expm1(torch.randn(n, *shape, generator=g) * 1.2 + 0.5).abs()
    u = torch.rand(n, *shape, generator=g)
    x[u < 0.01] = -x[u < 0.01]
    x[(u >= 0.01) & (u < 0.011)] = float('nan')
    x[(u >= 0.011) & (u < 0.0115)] = float('inf')
    x[(u >= 0.0115) & (u < 0.012)] = float('-inf')
    x[(u >= 0.012) & (u < 0.03)] = 0.0
    x[(u >= 0.03) & (u < 0.035)] = 5e-6     # below the valid threshold
    return x


@pytest.mark.parametrize('n,shape,black', [(7, (128, 128), 3.0), (40, (64, 96), 15.0), (3, (33, 17), 5.0), (2, (512, 512), 3.0)])
def test_robust_stats_equal_the_exact_population_percentiles(cuda, n, shape, black):
    """srb_robust_hist radix select == np.percentile over every valid pixel (oracle/aux.py robust_stats without the
    reference's random subsample), bit for bit; and the reference-style subsample estimate scatters around it."""
    from oracle import aux
    from superresolution_b200.normalization import robust_stats
    g = torch.Generator().manual_seed(100 + n)
    x = _sky(g, n, shape)
    ref_min, ref_max = aux.robust_stats([im.numpy() for im in x], black)
    chunks = [x[:n // 2].to(cuda), x[n // 2:].to(cuda)]          # uneven chunks, incl. sizes not divisible by 4
    got_min, got_max = robust_stats(chunks, black)
    assert got_min.dtype == np.float32 and got_max.dtype == np.float32
    assert got_min == np.float32(ref_min) and got_max == np.float32(ref_max), (got_min, ref_min, got_max, ref_max)
    got2 = robust_stats(lambda: iter([x.to(cuda).view(-1)]), black)  # streaming form, one flat batch
    assert got2 == (got_min, got_max)
    torch.manual_seed(0)
    est_min, est_max = aux.robust_stats([im.numpy() for im in x], black, num_samples_per_img=4000)
    assert abs(est_min - ref_min) <= 0.1 * (ref_max - ref_min)


def test_robust_stats_without_valid_pixels_falls_back(cuda):
    from superresolution_b200.normalization import robust_stats
    x = torch.full((4, 32, 32), float('nan'))
    x[0] = -3.0
    x[1] = 0.0
    assert robust_stats([x.to(cuda)]) == (np.float32(0.0), np.float32(1.0))


def test_tiff16_and_preview_glue(cuda, tmp_path):
    """save_as_tiff16 (Modello_5.py:29-33), the TIFF load of src/dataset.py:24-41 and the [LR|SR|HR] strip of
    Modello_5.py:83-85 against their reference expressions (PIL / torch on the CPU)."""
    from PIL import Image
    from superresolution_b200 import io16
    g = torch.Generator().manual_seed(21)
    sr = torch.rand(1, 1, 64, 96, generator=g) * 1.2 - 0.1           # values outside [0,1] exercise the clip
    io16.save_as_tiff16(sr.to(cuda), tmp_path / 'sr.tiff')
    want = (np.clip(sr.squeeze().numpy(), 0, 1) * 65535).astype(np.uint16)
    got = np.array(Image.open(tmp_path / 'sr.tiff'))
    assert got.dtype == np.uint16 and np.array_equal(got, want)
    back = io16.load_tiff_as_tensor(tmp_path / 'sr.tiff', cuda)
    ref = torch.from_numpy(np.array(Image.open(tmp_path / 'sr.tiff'), dtype=np.float32) / 65535.0).unsqueeze(0)
    assert back.dtype == torch.uint16 and torch.equal(back.cpu().float() / 65535.0, ref)
    for (h, w), (H, W) in (((16, 24), (64, 96)), ((20, 30), (64, 96))):  # integer and fractional nearest ratios
        lr = torch.rand(2, 1, h, w, generator=g)
        s2 = torch.rand(2, 1, H, W, generator=g) * 1.2 - 0.1
        hr = torch.rand(2, 1, H, W, generator=g)
        comp = torch.cat((torch.nn.functional.interpolate(lr, size=(H, W), mode='nearest'), s2, hr), dim=3).clamp(0, 1)
        out = io16.preview_strip(lr.to(cuda), s2.to(cuda), hr.to(cuda))
        assert torch.equal(out.cpu(), comp)
        q = io16.preview_strip(lr.to(cuda), s2.to(cuda), hr.to(cuda), quantize=True)
        assert torch.equal(q.cpu(), comp.mul(255).add_(0.5).clamp_(0, 255).to(torch.uint8))  # torchvision save_image
