"""train_step over two ranks (NCCL): the gradient all-reduce that replaces nn.DataParallel's reduce-add
(scripts/Modello_supporto.py:103-105, 125-141), launched in two parts — the stage-2 range as soon as the stage-2
micro-batches are done, under the stage-1 backward, then the stage-1 range.  Checks: the reduced flat gradient equals the
sum of the two ranks' local gradients, and both ranks hold bit-identical parameters after the step.
Needs >= 2 GPUs (run with `gpurun --gpus 2`); skipped otherwise."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.losses import CombinedLoss
    from superresolution_b200.optim import FusedAdamW
    from superresolution_b200.training import stage2_offset, train_step
    dev = torch.device('cuda', rank)
    torch.manual_seed(0)
    model = HybridSuperResolutionModel(output_size=256, smoothing='none').to(dev).train()
    opt = FusedAdamW(model.parameters(), lr=1e-3, weight_decay=0.0)
    crit = CombinedLoss()
    g = torch.Generator().manual_seed(50 + rank)
    x = torch.rand(2, 1, 64, 64, generator=g).to(dev)
    hr = (torch.rand(2, 1, 256, 256, generator=g) * 0.05).to(dev)
    cut = stage2_offset(model, opt)
    # local gradient of this rank (no exchange, no update: lr irrelevant because the parameters are restored)
    p0 = opt.flat_param.clone()
    train_step(model, opt, crit, x, hr, micro_batch=1, max_norm=1.0, world=1, drop_path_rate=0.0)
    local = opt.flat_grad.clone()
    opt.flat_param.copy_(p0)
    opt.exp_avg.zero_(); opt.exp_avg_sq.zero_(); opt._step = 0
    loss = train_step(model, opt, crit, x, hr, micro_batch=1, max_norm=1.0, world=world, drop_path_rate=0.0)
    torch.cuda.synchronize()
    torch.save({'local': local.cpu(), 'reduced': opt.flat_grad.cpu(), 'param': opt.flat_param.cpu(), 'cut': cut,
                'numel': opt.numel, 'loss': float(loss)}, os.path.join(out_dir, f'rank{rank}.pt'))
    dist.destroy_process_group()


def test_train_step_two_ranks_overlapped_allreduce(cuda, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r = [torch.load(os.path.join(tmp_path, f'rank{k}.pt')) for k in range(2)]
    assert 0 < r[0]['cut'] < r[0]['numel'] and r[0]['cut'] == r[1]['cut']  # stage 2 is the tail: something overlaps
    want = r[0]['local'] + r[1]['local']
    for k in range(2):
        # the attention-bias table gradients are summed with atomics (order varies between the two passes)
        err = (r[k]['reduced'] - want).abs().max().item()
        assert err <= 1e-5 * want.abs().max().item() + 1e-7, err
    assert torch.equal(r[0]['reduced'], r[1]['reduced'])
    assert torch.equal(r[0]['param'], r[1]['param'])
