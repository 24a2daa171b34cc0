"""BASELINE config 4: hybrid-model training step (RRDBNet + HAT), batch 32 of 128x128 patches per GPU, loss =
src/losses.py, smoothing='none' (scripts/Modello_supporto.py:101), gradient SUM all-reduce over the ranks (NCCL), fused
unscale + clip(1.0) + AdamW(5e-4, wd 0) — superresolution_b200.training.train_step.  The batch is taken in micro-batches
(default 16 patches, stage 2 only; stage 1 runs on the whole batch) whose gradients accumulate in the optimizer's flat
fp32 buffer: HAT's saved activations do not fit 180 GB at 32 patches.  Algorithmic work: 3 x 725.74 GFLOP per patch (forward + data gradient + weight gradient).

usage: python tools/bench_train.py [--batch 32] [--micro 4] [--steps 3] [--warmup 1] [--stage hybrid|stage1]
       python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/bench_train.py"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel  # noqa: E402
from superresolution_b200.losses import CombinedLoss  # noqa: E402
from superresolution_b200.optim import FusedAdamW, allreduce_gradients  # noqa: E402
from superresolution_b200.training import _head, _tail  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=32)
    ap.add_argument('--micro', type=int, default=16)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=1)
    a = ap.parse_args()
    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', 0)))
    dev = torch.device('cuda', torch.cuda.current_device())
    if world > 1:
        dist.init_process_group('nccl')
    torch.manual_seed(0)
    model = HybridSuperResolutionModel(smoothing='none').to(dev).train()
    opt = FusedAdamW(model.parameters(), lr=5e-4, weight_decay=0.0)
    crit = CombinedLoss()
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    x = torch.rand(a.batch, 1, 128, 128, device=dev, generator=g)
    t = torch.rand(a.batch, 1, 512, 512, device=dev, generator=g) * 0.02
    t[torch.rand(t.shape, device=dev, generator=g) < 0.02] = 0.9  # ~2 % bright "stars": both loss branches fire

    ev = lambda: torch.cuda.Event(enable_timing=True)
    phases = {'stage1 forward': 0.0, 'stage2 forward+loss': 0.0, 'stage2 backward': 0.0, 'stage1 backward': 0.0,
              'allreduce': 0.0, 'optimizer': 0.0}

    def step(record):
        """superresolution_b200.training.train_step, unrolled here to put events around its phases"""
        opt.zero_grad()
        e0, e1 = ev(), ev()
        e0.record()
        y1 = _head(model, x, True)                      # stage 1 forward, whole batch
        e1.record()
        g1 = torch.empty_like(y1)
        marks = []
        for s in range(0, a.batch, a.micro):
            e = [ev() for _ in range(3)]
            e[0].record()
            ys = y1[s:s + a.micro].detach().requires_grad_(True)
            loss, _ = crit(_tail(model, ys, 0.1), t[s:s + a.micro])
            e[1].record()
            (loss * (ys.shape[0] / a.batch)).backward()
            g1[s:s + a.micro] = ys.grad
            e[2].record()
            marks.append(e)
        e2, e3, e4, e5 = ev(), ev(), ev(), ev()
        e2.record()
        y1.backward(g1)                                 # stage 1 backward, whole batch
        e3.record()
        if world > 1:
            allreduce_gradients(opt.flat_grad)
        e4.record()
        opt.step(grad_scale=1.0, max_norm=1.0, world=world)
        e5.record()
        if record:
            torch.cuda.synchronize()
            phases['stage1 forward'] += e0.elapsed_time(e1)
            for e in marks:
                phases['stage2 forward+loss'] += e[0].elapsed_time(e[1])
                phases['stage2 backward'] += e[1].elapsed_time(e[2])
            phases['stage1 backward'] += e2.elapsed_time(e3)
            phases['allreduce'] += e3.elapsed_time(e4)
            phases['optimizer'] += e4.elapsed_time(e5)
        return loss

    for _ in range(a.warmup):
        step(False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(a.steps):
        loss = step(True)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        per = ms.item() / a.steps
        flop = 3 * 725.74e9 * a.batch
        print(json.dumps({'workload': f'config 4: hybrid (RRDBNet + HAT) training step, batch {a.batch}/GPU of 128x128 in '
                                      f'stage-2 micro-batches of {a.micro}, weighted-L1 loss, grad all-reduce, clip 1.0, AdamW',
                          'n_gpus': world, 'ms_per_step': round(per, 1),
                          'samples_per_s': round(world * a.batch / (per / 1e3), 2),
                          'model_tflops_per_gpu': round(flop / (per / 1e3) / 1e12, 1),
                          'phase_ms': {k: round(v / a.steps, 1) for k, v in phases.items()},
                          'loss': round(loss.item(), 5),
                          'peak_mem_gb': round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
