"""Config-4 training step over the ranks of a torchrun launch with the gradient exchange (a) after the whole backward,
(b) split: stage-2 range under the stage-1 backward.  usage: torchrun --nproc-per-node N tools/bench_allreduce_modes.py"""
import os, sys, json, torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200 import ops
from superresolution_b200.architecture import HybridSuperResolutionModel
from superresolution_b200.losses import CombinedLoss
from superresolution_b200.optim import FusedAdamW, allreduce_gradients_start, allreduce_gradients_wait
from superresolution_b200.training import _head, _tail, stage2_offset

rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
dev = torch.device('cuda', int(os.environ['LOCAL_RANK']))
torch.cuda.set_device(dev)
dist.init_process_group('nccl', device_id=dev)
torch.manual_seed(0)
model = HybridSuperResolutionModel(smoothing='none').to(dev).train()
opt = FusedAdamW(model.parameters(), lr=5e-4, weight_decay=0.0)
crit = CombinedLoss()
cut = stage2_offset(model, opt)
g = torch.Generator(device=dev).manual_seed(10 + rank)
batch, micro = 32, 16
x = torch.rand(batch, 1, 128, 128, device=dev, generator=g)
t = torch.rand(batch, 1, 512, 512, device=dev, generator=g) * 0.02


def step(overlap):
    opt.zero_grad()
    y1 = _head(model, x, True)
    g1 = torch.empty_like(y1)
    with ops.weight_pack_cache():
        for s in range(0, batch, micro):
            ys = y1[s:s + micro].detach().requires_grad_(True)
            loss, _ = crit(_tail(model, ys, 0.1), t[s:s + micro])
            (loss * (ys.shape[0] / batch)).backward()
            g1[s:s + micro] = ys.grad
    works = allreduce_gradients_start(opt.flat_grad, lo=cut) if overlap else []
    y1.backward(g1)
    works += allreduce_gradients_start(opt.flat_grad, hi=cut if overlap else None)
    allreduce_gradients_wait(works)
    opt.step(grad_scale=1.0, max_norm=1.0, world=world)


res = {}
for name, ov in (('after_backward', False), ('stage2_under_stage1_backward', True), ('after_backward_again', False)):
    step(ov)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        step(ov)
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 3], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    res[name] = round(ms.item(), 1)
if rank == 0:
    print(json.dumps({'world': world, 'nccl_max_ctas': os.environ.get('NCCL_MAX_CTAS'), 'ms_per_step': res}))
dist.destroy_process_group()
