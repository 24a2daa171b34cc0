"""Whole-model training path (SURVEY.md §8 a2 + a12): superresolution_b200.training.hybrid_forward / train_step and the
operators that join the two stages (AntiCheckerboardLayer blur, bicubic resize branch of src/architecture.py:80-81) against
the fp32 oracle (oracle/hybrid.py = the reference's forward over the re-derived RRDBNet / HAT) and PyTorch autograd.

Bars: blur and bicubic are fp32 operators -> 1e-5 absolute against ATen.  The hybrid gradients run through ~400 bf16
layers whose LeakyReLU sign patterns differ from an fp32 forward at ~0.05-0.3 % of the elements (see
test_training_gpu.py / test_training_hat_gpu.py, where each stage is checked against a sign-matched reference to 2-4 %):
here the composition is checked, so the bars are per-stage medians of the relative L2 error and the cosine of the whole
flattened gradient."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize('mode', ['light', 'balanced', 'strong'])
def test_blur_abi_matches_reference_layer_and_is_self_adjoint(cuda, mode):
    from oracle.hybrid import AntiCheckerboardLayer as RefBlur
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(11)
    x = torch.rand(3, 1, 40, 72, generator=g)
    ref = RefBlur(mode)(x)
    got = ops.blur(x.to(cuda), mode)
    assert (got.cpu() - ref).abs().max().item() <= 1e-6
    # <blur(x), r> == <x, blur(r)>: the backward used by the training path
    r = torch.rand(3, 1, 40, 72, generator=g)
    lhs = (got.double() * r.to(cuda).double()).sum().item()
    rhs = (x.to(cuda).double() * ops.blur(r.to(cuda), mode).double()).sum().item()
    assert abs(lhs - rhs) <= 1e-6 * abs(lhs)
    u16 = ops.blur(x.to(cuda), mode, out_dtype=torch.uint16)
    exp = (ref.clamp(0, 1) * 65535.0).to(torch.int32)
    assert (u16.cpu().to(torch.int32) - exp).abs().max().item() <= 1


@pytest.mark.parametrize('hw_in,hw_out', [((128, 256), (192, 192)), ((64, 64), (512, 512)), ((96, 128), (37, 53))])
def test_bicubic_resize_forward_backward_match_aten(cuda, hw_in, hw_out):
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(12)
    x = torch.rand(2, 1, *hw_in, generator=g, requires_grad=True)
    ref = F.interpolate(x, size=hw_out, mode='bicubic', align_corners=False)
    r = torch.randn(ref.shape, generator=g)
    (ref * r).sum().backward()
    got = ops.bicubic_resize(x.detach().to(cuda), *hw_out)
    assert got.shape == ref.shape
    assert (got.cpu() - ref.detach()).abs().max().item() <= 1e-5
    gx = ops.bicubic_resize_bwd(r.to(cuda), *hw_in)
    assert (gx.cpu() - x.grad).abs().max().item() <= 1e-4 * max(1.0, x.grad.abs().max().item())


@pytest.mark.parametrize('hw_in,hw_out', [((128, 256), (48, 96)), ((512, 512), (192, 192)), ((64, 64), (256, 256))])
def test_antialiased_bicubic_resize_matches_aten(cuda, hw_in, hw_out):
    """F.interpolate(..., mode='bicubic', antialias=True) of the stale Backup/src/architecture.py:91-92."""
    from superresolution_b200 import ops
    g = torch.Generator().manual_seed(13)
    x = torch.rand(2, 1, *hw_in, generator=g)
    ref = F.interpolate(x, size=hw_out, mode='bicubic', align_corners=False, antialias=True)
    got = ops.bicubic_resize(x.to(cuda), *hw_out, antialias=True)
    assert got.shape == ref.shape
    assert (got.cpu() - ref).abs().max().item() <= 2e-5


def test_inference_resize_branch_matches_oracle(cuda):
    """output_size != 4W: stage1 -> s1 -> stage2 -> s2 -> bicubic -> sf (src/architecture.py:74-82), bf16 bar 50 dB."""
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    torch.manual_seed(0)
    o = Oracle(output_size=96, smoothing='balanced').eval()
    m = Model(output_size=96, smoothing='balanced')
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda).eval()
    g = torch.Generator().manual_seed(1)
    x = torch.randint(0, 65536, (1, 1, 32, 32), generator=g).float() / 65535.0
    with torch.no_grad():
        ref = o(x)
        got = m(x.to(cuda)).cpu()
    assert got.shape == ref.shape == (1, 1, 96, 96)
    mse = (got.double() - ref.double()).pow(2).mean().item()
    assert 10 * math.log10(1.0 / max(mse, 1e-20)) >= 50.0


@pytest.mark.parametrize('smoothing,output_size', [('none', 256), ('balanced', 192)])
def test_hybrid_forward_backward_matches_oracle_autograd(cuda, smoothing, output_size):
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    from superresolution_b200.training import hybrid_forward
    torch.manual_seed(2)
    o = Oracle(output_size=output_size, smoothing=smoothing).eval()   # eval: DropPath off on both sides
    m = Model(output_size=output_size, smoothing=smoothing)
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda).train()
    g = torch.Generator().manual_seed(8)
    x = torch.randint(0, 65536, (1, 1, 32, 64), generator=g).float() / 65535.0
    yo = o(x)
    R = torch.randn(yo.shape, generator=g)
    (yo * R).sum().backward()
    y = hybrid_forward(m, x.to(cuda), drop_path_rate=0.0)
    assert y.shape == yo.shape and y.requires_grad
    mse = (y.detach().cpu().double() - yo.detach().double()).pow(2).mean().item()
    psnr = 10 * math.log10(1.0 / max(mse, 1e-20))
    print(f'forward psnr vs fp32 oracle {psnr:.2f} dB')
    assert psnr >= 50.0
    (y * R.to(cuda)).sum().backward()
    ref = dict(o.named_parameters())
    # plain fp32 oracle, LeakyReLU sign flips included (HAT's LeakyReLU(0.01): a flipped element changes its gradient
    # 100x): measured stage 2 median 2.6 / 7.3 %, p90 3.2 / 9.4 %, cosine 0.9998 / 0.9983; stage 1 median 8.0 / 11.6 %,
    # p90 10.8 / 15.0 %, cosine 0.998 / 0.995 (balanced-192 / none-256).  Bars ~1.6-2x the worse case; the sign-matched
    # test below is the tight one.
    for stage, med_bar, p90_bar, cos_bar in (('stage2.', 0.12, 0.16, 0.997), ('stage1.', 0.20, 0.28, 0.99)):
        errs, a, b = [], [], []
        for n, p in m.named_parameters():
            if not n.startswith(stage):
                continue
            assert p.grad is not None and p.grad.shape == ref[n].grad.shape, n
            errs.append(_rel(p.grad.cpu(), ref[n].grad))
            a.append(p.grad.cpu().double().flatten())
            b.append(ref[n].grad.double().flatten())
        a, b = torch.cat(a), torch.cat(b)
        cos = (a @ b / (a.norm() * b.norm())).item()
        errs.sort()
        print(f'{stage} {len(errs)} tensors: median rel err {errs[len(errs) // 2]:.4f}  p90 {errs[int(len(errs) * .9)]:.4f} '
              f'max {errs[-1]:.4f}  cosine {cos:.5f}  |g| {a.norm():.4g} vs {b.norm():.4g}')
        assert errs[len(errs) // 2] <= med_bar and errs[int(len(errs) * .9)] <= p90_bar
        assert cos >= cos_bar


def _grad_stats(tag, named_got, ref):
    """per-tensor relative L2 error and cosine of every gradient tensor; cosines only for tensors whose reference norm
    is above 1e-3 of the largest one (below that floor bf16 storage noise dominates a near-zero gradient)."""
    errs, coss, norms = {}, {}, {}
    for n, p in named_got:
        assert p.grad is not None and p.grad.shape == ref[n].grad.shape, n
        a, b = p.grad.cpu().double().flatten(), ref[n].grad.double().flatten()
        errs[n] = ((a - b).norm() / b.norm().clamp_min(1e-30)).item()
        coss[n] = (a @ b / (a.norm() * b.norm()).clamp_min(1e-30)).item()
        norms[n] = b.norm().item()
    floor = 1e-3 * max(norms.values())
    v = sorted(errs.values())
    live = {n: c for n, c in coss.items() if norms[n] >= floor}
    worst = max(errs, key=errs.get)
    low = min(live, key=live.get)
    st = {'median': v[len(v) // 2], 'p90': v[int(len(v) * .9)], 'max': v[-1], 'min_cos': live[low],
          'max_live': max(errs[n] for n in live)}
    print(f'{tag}: {len(v)} tensors  median {st["median"]:.4f}  p90 {st["p90"]:.4f}  max {st["max"]:.4f} ({worst}, '
          f'|g| {norms[worst]:.3g}; above the floor {st["max_live"]:.4f})  min cosine over {len(live)} tensors above the floor {st["min_cos"]:.5f} ({low})')
    return st


@pytest.mark.parametrize('drop,smoothing,output_size', [(False, 'balanced', 192), (True, 'none', 256)])
def test_hybrid_gradients_match_rounding_matched_oracle(cuda, drop, smoothing, output_size, monkeypatch):
    """the composition stage1 -> s1 -> stage2 -> s2 -> (resize) -> sf, forward and backward,
    against autograd over the oracle with the kernels' rounding points in stage 1 (tests/_emulate.py) and the sign
    pattern of HAT's one LeakyReLU taken from the libsrb forward — so the bars can sit at ~2x the measured errors instead
    of absorbing sign flips.  drop=True: train mode with the SAME DropPath factors on both sides (rate 0.4 so that
    branches really drop).  Bars (measured on B200, printed): see the asserts."""
    import oracle.hat as ohat
    from _emulate import _emulated_rrdbnet
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    from superresolution_b200.training import hybrid_forward
    from superresolution_b200.training_hat import drop_path_masks
    torch.manual_seed(2)
    o = Oracle(output_size=output_size, smoothing=smoothing)
    m = Model(output_size=output_size, smoothing=smoothing)
    m.load_state_dict(o.state_dict(), strict=True)
    m = m.to(cuda).train()
    B = 2 if drop else 1
    g = torch.Generator().manual_seed(8)
    x = torch.randint(0, 65536, (B, 1, 32, 64), generator=g).float() / 65535.0
    masks = None
    if drop:
        o.train()
        torch.manual_seed(3)
        masks = [(None, None)] + [(drop_path_masks(B, 0.4, True, 'cpu'), drop_path_masks(B, 0.4, True, 'cpu'))
                                  for _ in range(23)]
        assert any(v == 0.0 for mm in masks[1:] for half in mm for v in half)
        it = iter([half for mm in masks[1:] for half in mm])

        def shared_drop_path(t, p=0.0, training=False):
            return t * torch.as_tensor(next(it), dtype=t.dtype).view(-1, *([1] * (t.ndim - 1)))

        monkeypatch.setattr(ohat, 'drop_path', shared_drop_path)
    else:
        o.eval()
    dbg = {}
    y = hybrid_forward(m, x.to(cuda), drop_path_rate=0.4 if drop else 0.0, masks=masks, debug=dbg)
    pos = (dbg['u'].float() > 0).permute(0, 3, 1, 2).cpu()

    class SignMatchedLeakyReLU(torch.nn.Module):
        def forward(self, t):
            return t * torch.where(pos, 1.0, 0.01)

    o.stage2.conv_before_upsample[1] = SignMatchedLeakyReLU()
    yo = o.stage2(o.s1(_emulated_rrdbnet(o.stage1, x)))
    yo = o.s2(yo)
    if yo.shape[-1] != output_size:
        yo = F.interpolate(yo, size=(output_size, output_size), mode='bicubic', align_corners=False)
    yo = o.sf(yo)
    assert y.shape == yo.shape and y.requires_grad
    fe = _rel(y.detach().cpu() - y.detach().cpu().mean(), yo.detach() - yo.detach().mean())
    R = torch.randn(yo.shape, generator=g)
    (yo * R).sum().backward()
    (y * R.to(cuda)).sum().backward()
    ref = dict(o.named_parameters())
    s2 = _grad_stats('stage2', [(n, p) for n, p in m.named_parameters() if n.startswith('stage2.')], ref)
    s1 = _grad_stats('stage1', [(n, p) for n, p in m.named_parameters() if n.startswith('stage1.')], ref)
    print(f'forward rel err (mean removed) {fe:.4f}')
    assert fe <= 0.03
    # measured (B200): stage 2 median 0.7 / 1.1 %, p90 1.0 / 1.5 %, min cosine 0.9999; stage 1 median 2.4 / 3.1 %, p90 3.1 / 3.6 %,
    # max 8.0 / 5.4 %, min cosine 0.9990 / 0.9995 (eval / shared-DropPath train mode).  Bars ~2x.  `max` over ALL tensors is
    # dominated by near-zero gradients (|g| ~1e-5 of the largest: channel-attention 1x1 weights) and only sanity-bounded.
    assert s2['median'] <= 0.025 and s2['p90'] <= 0.035 and s2['max_live'] <= 0.06 and s2['min_cos'] >= 0.999
    assert s1['median'] <= 0.06 and s1['p90'] <= 0.075 and s1['max_live'] <= 0.16 and s1['min_cos'] >= 0.997
    # (the stage-2 maximum sits on a channel-attention 1x1 weight with |g| ~6e-6, five orders below the largest tensors:
    # bf16 noise of everything upstream decides its direction; 0.3 - 1.4 across kernel revisions)
    assert s2['max'] <= 2.5 and s1['max'] <= 0.2


def test_train_step_micro_batches_equal_full_batch_and_reduce_loss(cuda):
    """train_step = scripts/Modello_supporto.py:125-141: the gradient of a batch taken in micro-batches equals the
    gradient of the batch taken at once (deterministic kernels: bit-equal is not required, 1e-3 relative is), and a few
    steps lower the reference loss."""
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    from superresolution_b200.losses import CombinedLoss
    from superresolution_b200.optim import FusedAdamW
    from superresolution_b200.training import hybrid_forward, train_step
    torch.manual_seed(3)
    m = Model(output_size=128, smoothing='none').to(cuda).train()
    opt = FusedAdamW(m.parameters(), lr=2e-4, weight_decay=0.0)
    crit = CombinedLoss()
    g = torch.Generator().manual_seed(9)
    x = (torch.randint(0, 65536, (2, 1, 32, 32 * 2), generator=g).float() / 65535.0).to(cuda)
    t = F.interpolate(x, scale_factor=4, mode='bilinear', align_corners=False)
    t[:, :, ::17, ::13] = 0.9  # bright "stars": the 500x branch of src/losses.py fires
    m.output_size = t.shape[-1]
    # (a) the whole batch through hybrid_forward at once; (b) train_step's plan (stage 1 on the batch, stage 2 in slices of
    # one patch, input gradients collected) with a zero learning rate so that the parameters stay put
    opt.zero_grad()
    loss, _ = crit(hybrid_forward(m, x, accumulate_into_grad=True, drop_path_rate=0.0), t)
    loss.backward()
    g_full, l_full = opt.flat_grad.clone(), loss.item()
    opt.param_groups[0]['lr'] = 0.0
    l_step = train_step(m, opt, crit, x, t, micro_batch=1, max_norm=1.0, drop_path_rate=0.0).item()
    opt.param_groups[0]['lr'] = 2e-4
    err = _rel(opt.flat_grad, g_full)
    print(f'train_step (micro-batch 1) vs full-batch gradient rel err {err:.2e}, loss {l_step:.6f} vs {l_full:.6f}')
    assert err <= 1e-3 and abs(l_step - l_full) <= 1e-4 * abs(l_full)
    losses = [train_step(m, opt, crit, x, t, micro_batch=1, max_norm=1.0).item() for _ in range(5)]
    print('losses', [round(v, 4) for v in losses])
    assert all(math.isfinite(v) for v in losses) and losses[-1] < losses[0]


def test_reference_training_loop_runs_unchanged(cuda):
    """The loop body of scripts/Modello_supporto.py:125-141, verbatim (autocast, GradScaler, unscale_, clip_grad_norm_,
    stock optim.AdamW, zero_grad), over the drop-in module and loss: forward and backward run on libsrb, parameters move,
    the loss falls, and a state dict saved afterwards loads back strictly (:177-185)."""
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.losses import CombinedLoss
    torch.manual_seed(4)
    device = cuda
    model = HybridSuperResolutionModel(output_size=256, smoothing='none', device=device).to(device)
    optimizer = torch.optim.AdamW(model.parameters(), lr=2e-4, weight_decay=0)
    criterion = CombinedLoss().to(device)
    scaler = torch.cuda.amp.GradScaler()
    ACCUM_STEPS = 1
    g = torch.Generator().manual_seed(10)
    lr_img = torch.randint(0, 65536, (2, 1, 64, 64), generator=g).float() / 65535.0
    hr_img = F.interpolate(lr_img, scale_factor=4, mode='bilinear', align_corners=False)
    batch = {'lr': lr_img, 'hr': hr_img}
    before = {k: v.detach().clone() for k, v in model.state_dict().items()}
    model.train()
    optimizer.zero_grad()
    losses = []
    for i in range(4):
        lr = batch['lr'].to(device, non_blocking=True)
        hr = batch['hr'].to(device, non_blocking=True)
        with torch.cuda.amp.autocast():
            pred = model(lr)
            loss, _ = criterion(pred, hr)
            loss = loss / ACCUM_STEPS
        scaler.scale(loss).backward()
        if (i + 1) % ACCUM_STEPS == 0:
            scaler.unscale_(optimizer)
            gn = torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            scaler.step(optimizer)
            scaler.update()
            optimizer.zero_grad()
        assert math.isfinite(gn.item()) and gn.item() > 0
        losses.append(loss.item() * ACCUM_STEPS)
    print('losses', [round(v, 4) for v in losses])
    assert losses[-1] < losses[0]
    after = model.state_dict()
    moved = sum(int(not torch.equal(before[k], after[k])) for k in before if before[k].is_floating_point())
    assert moved >= 1200  # every weight / bias tensor of both stages received a gradient and a step
    model.eval()
    with torch.inference_mode(), torch.cuda.amp.autocast():
        v_pred = model(batch['lr'].to(device))          # validation forward of :159-166: the fused inference path
    assert v_pred.shape == (2, 1, 256, 256) and torch.isfinite(v_pred).all()
    fresh = HybridSuperResolutionModel(output_size=256, smoothing='none')
    fresh.load_state_dict({k: v.cpu() for k, v in after.items()}, strict=True)
