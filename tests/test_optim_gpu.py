"""Optimizer half of the training step: libsrb FusedAdamW against torch.optim.AdamW +
GradScaler.unscale_ + clip_grad_norm_ (the sequence of reference scripts/Modello_supporto.py:133-141) on the same
seeded parameters and gradients.  fp32 arithmetic: tolerance 2e-6 absolute on O(1) parameters."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(64, 4, 3, 3), (64,), (32, 64, 3, 3), (96, 96), (3,), (225, 6), (1, 64, 3, 3), (5,)]


def _make(cuda, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(s, generator=g).to(cuda)) for s in SHAPES]


def _grads(seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(s, generator=g) * scale for s in SHAPES]


def test_plain_step_matches_torch_adamw(cuda):
    from superresolution_b200.optim import FusedAdamW
    a, b = _make(cuda, 0), _make(cuda, 0)
    ref = torch.optim.AdamW(a, lr=5e-4, weight_decay=0.01)
    opt = FusedAdamW(b, lr=5e-4, weight_decay=0.01)
    for it in range(6):
        for p, q, g in zip(a, b, _grads(10 + it)):
            p.grad = g.to(cuda)
            q.grad.copy_(g.to(cuda))
        ref.step()
        opt.step()
        ref.zero_grad()
        opt.zero_grad()
    for p, q in zip(a, b):
        assert torch.allclose(p, q, rtol=0, atol=2e-6), (p - q).abs().max().item()
        assert q.grad is not None and q.grad.abs().max().item() == 0.0  # views survive zero_grad


def test_step_survives_dropped_gradient_views(cuda):
    """after ``p.grad = None`` (what ``model.zero_grad()`` does by default) autograd allocates fresh .grad
    tensors that no longer alias the flat buffer: the step must use THEM, not the stale buffer, in the plain and in the
    fused form; a parameter left without a gradient contributes a zero gradient."""
    from superresolution_b200.optim import FusedAdamW
    for fused in (False, True):
        a, b = _make(cuda, 1), _make(cuda, 1)
        ref = torch.optim.AdamW(a, lr=5e-4, weight_decay=0.0)
        opt = FusedAdamW(b, lr=5e-4, weight_decay=0.0)
        for it in range(4):
            for k, (p, q, g) in enumerate(zip(a, b, _grads(30 + it))):
                q.grad = None                      # drop the view ...
                if k == 2 and it == 1:
                    p.grad = torch.zeros_like(p)   # ... and leave one parameter without a gradient for one step
                    continue
                p.grad = g.to(cuda)
                q.grad = g.to(cuda).clone()        # ... a fresh tensor, as autograd would allocate
            ref.step()
            if fused:
                opt.step(grad_scale=1.0, max_norm=1e9, world=1)
            else:
                opt.step()
        for p, q in zip(a, b):
            assert torch.allclose(p, q, rtol=0, atol=2e-6), (fused, (p - q).abs().max().item())
            assert q.grad.data_ptr() >= opt.flat_grad.data_ptr()  # re-pointed into the flat buffer


def test_fused_unscale_clip_step_matches_reference_sequence(cuda):
    from superresolution_b200.optim import FusedAdamW
    a, b = _make(cuda, 1), _make(cuda, 1)
    ref = torch.optim.AdamW(a, lr=5e-4, weight_decay=0.0)
    opt = FusedAdamW(b, lr=5e-4, weight_decay=0.0, bf16_shadow=True)
    scale = 1024.0
    for it in range(4):
        gs = _grads(20 + it, scale=scale * (3.0 if it % 2 else 0.01))  # clipped and unclipped steps
        for p, q, g in zip(a, b, gs):
            p.grad = g.to(cuda) / scale  # what scaler.unscale_ leaves
            q.grad.copy_(g.to(cuda))
        norm_ref = torch.nn.utils.clip_grad_norm_(a, 1.0)
        norm = opt.grad_norm(grad_scale=scale)
        assert abs(norm.item() - norm_ref.item()) <= 1e-5 * max(1.0, norm_ref.item())
        ref.step()
        opt.step(grad_scale=scale, max_norm=1.0)
        assert opt.found_inf().item() == 0.0
    for p, q in zip(a, b):
        assert torch.allclose(p, q, rtol=0, atol=2e-6), (p - q).abs().max().item()
    assert torch.equal(opt.shadow.float(), opt.flat_param.bfloat16().float())


def test_nonfinite_gradient_skips_the_step(cuda):
    from superresolution_b200.optim import FusedAdamW
    b = _make(cuda, 2)
    opt = FusedAdamW(b, lr=1e-3, weight_decay=0.0)
    before = opt.flat_param.clone()
    for q, g in zip(b, _grads(30)):
        q.grad.copy_(g.to(cuda))
    b[3].grad[5, 7] = float('inf')
    opt.step(grad_scale=65536.0, max_norm=1.0)
    assert opt.found_inf().item() == 1.0
    assert torch.equal(opt.flat_param, before) and opt.exp_avg.abs().max().item() == 0.0
    b[3].grad[5, 7] = float('nan')
    opt.step(grad_scale=65536.0, max_norm=1.0)
    assert opt.found_inf().item() == 1.0 and torch.equal(opt.flat_param, before)


def test_drives_model_parameters_and_repack(cuda):
    """The optimizer re-homes the model's parameters into its flat buffer; state_dict keys/values are unchanged and
    the model's packed-weight cache notices the update (version counters are bumped)."""
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.losses import CombinedLoss
    from superresolution_b200.optim import FusedAdamW
    torch.manual_seed(0)
    m = HybridSuperResolutionModel(output_size=128, smoothing='none').to(cuda).eval()
    sd0 = {k: v.clone() for k, v in m.state_dict().items()}
    opt = FusedAdamW(m.parameters(), lr=5e-4, weight_decay=0.0)
    assert list(m.state_dict().keys()) == list(sd0.keys())
    assert all(torch.equal(v, sd0[k]) for k, v in m.state_dict().items())
    x = torch.rand(1, 1, 32, 32, device=cuda)
    with torch.no_grad():
        y0 = m(x)
    # gradient on one tensor only (this test is about the optimizer): d loss / d conv_last.bias through the loss kernel
    t = torch.rand_like(y0)
    bias = m.stage2.conv_last.bias
    loss, _ = CombinedLoss()(y0 + bias.view(1, 1, 1, 1) * 0 + (bias - bias.detach()).view(1, 1, 1, 1), t)
    loss.backward()
    assert bias.grad.abs().sum().item() > 0
    opt.step(grad_scale=1.0, max_norm=1.0)
    with torch.no_grad():
        y1 = m(x)
    assert not torch.equal(y0, y1)  # repacked with the updated bias
    assert (m.stage2.conv_last.bias - sd0['stage2.conv_last.bias']).abs().max().item() > 0


def test_stage2_gradient_range_is_the_tail_of_the_flat_buffer(cuda):
    """train_step all-reduces flat_grad[cut:] (stage 2) under the stage-1 backward: the cut must separate the stages."""
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.optim import FusedAdamW
    from superresolution_b200.training import stage2_offset
    m = HybridSuperResolutionModel(smoothing='none').to(cuda)
    opt = FusedAdamW(m.parameters(), lr=1e-3)
    cut = stage2_offset(m, opt)
    base = opt.flat_grad.data_ptr()
    assert 0 < cut < opt.numel
    for n, p in m.named_parameters():
        off = (p.grad.data_ptr() - base) // 4
        assert (off >= cut) == n.startswith('stage2.'), n
    # a parameter list that interleaves the stages cannot be split: nothing is reduced early
    mixed = list(m.stage2.parameters())[:3] + list(m.stage1.parameters())[:3]
    opt2 = FusedAdamW(mixed, lr=1e-3)
    assert stage2_offset(m, opt2) == opt2.numel


def test_packed_weight_cache_lives_only_inside_its_scope(cuda):
    """train_step packs every stage-2 weight once per step (ops.weight_pack_cache); outside a scope nothing is cached,
    so a parameter edited between two forwards never meets a stale packed image."""
    from superresolution_b200 import ops
    p = torch.nn.Parameter(torch.randn(96, 32, 3, 3, device=cuda))
    w = p.detach()
    a = ops.packed_weight(p, w)
    with torch.no_grad():
        p.data.mul_(2.0)  # an edit torch's version counter does not see
    b = ops.packed_weight(p, w)
    assert a.data_ptr() != b.data_ptr() and torch.equal(b.float(), (2 * a.float()).bfloat16().float())
    with ops.weight_pack_cache():
        c1, d1 = ops.packed_weight(p, w), ops.packed_weight(p, w, dgrad=True)
        c2, d2 = ops.packed_weight(p, w), ops.packed_weight(p, w, dgrad=True)
        assert c1 is c2 and d1 is d2 and torch.equal(c1, b)
        assert torch.equal(d1, ops.pack_conv_weight_dgrad(w))
    with ops.weight_pack_cache():
        assert ops.packed_weight(p, w) is not c1  # a new scope (= a new step) packs again


def test_state_dict_round_trip_and_skipped_step_count(cuda):
    """state_dict carries step + both moments (a resumed run continues like torch.optim.AdamW would); a skipped fused step
    can be taken back out of the bias-correction count."""
    from superresolution_b200.optim import FusedAdamW
    a, b = _make(cuda, 4), _make(cuda, 4)
    oa, ob = FusedAdamW(a, lr=2e-3, weight_decay=0.01), FusedAdamW(b, lr=1e-4, weight_decay=0.0)
    for step in range(3):
        for q, g in zip(a, _grads(40 + step)):
            q.grad.copy_(g.to(cuda))
        oa.step()
    sd = oa.state_dict()
    assert sd['state'][0]['step'] == 3 and sd['param_groups'][0]['lr'] == 2e-3
    ob.load_state_dict(sd)
    ob.flat_param.copy_(oa.flat_param)
    for o, ps in ((oa, a), (ob, b)):
        for q, g in zip(ps, _grads(50)):
            q.grad.copy_(g.to(cuda))
        o.step()
    assert torch.equal(oa.flat_param, ob.flat_param) and torch.equal(oa.exp_avg_sq, ob.exp_avg_sq)
    # skipped step: the count is taken back, so the next update uses the bias correction of step 5, not 6
    a[0].grad[0, 0] = float('inf')
    oa.step(grad_scale=1.0, max_norm=1.0)
    assert oa.settle_skipped_step() is True and oa._step == 4
    assert oa.settle_skipped_step() is False and oa._step == 4  # settled once
