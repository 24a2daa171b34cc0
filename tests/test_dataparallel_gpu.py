"""nn.DataParallel is the reference's only multi-GPU mode (scripts/Modello_supporto.py:103-105; unwrap on save
:177-185).  Single process, one thread per device: exercises the per-device dynamic-shared-memory opt-in and SM count
of libsrb, the replica parameter lookup (replicate() moves parameters out of ``_parameters``) and the per-device
packed-weight / workspace caches.  Needs >= 2 GPUs (run with `gpurun --gpus 2`); skipped otherwise."""
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def two_gpus(cuda):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    return cuda


def _input(B, H, W, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 65536, (B, 1, H, W), generator=g).float() / 65535.0


def test_dataparallel_inference_matches_single_device(two_gpus):
    from superresolution_b200.architecture import HybridSuperResolutionModel
    torch.manual_seed(0)
    m = HybridSuperResolutionModel(output_size=256, smoothing='balanced').to('cuda:0').eval()
    x = _input(4, 64, 64, seed=5).to('cuda:0')
    with torch.no_grad():
        single = m(x)
        dp = nn.DataParallel(m, device_ids=[0, 1])
        par = dp(x)
        par2 = dp(x)  # second pass: cached per-device workspaces
    assert par.device == single.device and par.shape == single.shape == (4, 1, 256, 256)
    assert torch.equal(par, single) and torch.equal(par2, single)
    assert dp.module is m and 'stage1.conv_first.weight' in dp.module.state_dict()  # the save path of :177-185


def test_dataparallel_training_step_gradients(two_gpus):
    """forward + backward through DataParallel (train mode, smoothing='none' like Modello_supporto.py:101): the
    gradients reduced onto device 0 equal the single-device gradients of the same batch (bf16 kernels: the sum over
    the two half-batches is taken in a different order, hence a tolerance)."""
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.losses import CombinedLoss
    torch.manual_seed(0)
    m = HybridSuperResolutionModel(output_size=256, smoothing='none').to('cuda:0').train()
    crit = CombinedLoss()
    x = _input(2, 64, 64, seed=6).to('cuda:0')
    hr = _input(2, 256, 256, seed=7).to('cuda:0') * 0.05

    def grads(model):
        for p in m.parameters():
            p.grad = None
        torch.manual_seed(123)  # DropPath draws
        loss, _ = crit(model(x), hr)
        loss.backward()
        return loss.detach().clone(), {n: p.grad.detach().clone() for n, p in m.named_parameters() if p.grad is not None}

    # DropPath masks are drawn per replica thread from different generators -> compare with stochastic depth off
    import superresolution_b200.training as T
    orig = T.hybrid_forward
    T.hybrid_forward = lambda model, xx, accumulate_into_grad=False, drop_path_rate=0.0: orig(model, xx, accumulate_into_grad, 0.0)
    try:
        l1, g1 = grads(m)
        lp, gp = grads(nn.DataParallel(m, device_ids=[0, 1]))
    finally:
        T.hybrid_forward = orig
    assert set(g1) == set(gp) and len(g1) > 1200
    assert abs(l1.item() - lp.item()) <= 1e-3 * abs(l1.item()) + 1e-6
    worst = 0.0
    for n in g1:
        a, b = g1[n].double(), gp[n].double()
        if a.norm() > 1e-8:
            worst = max(worst, ((a - b).norm() / a.norm()).item())
    assert worst <= 0.1, worst
