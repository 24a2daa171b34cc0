"""Training-step measurement of the STAGE-1 (RRDBNet) half of the hybrid model alone: batch 32 of 128x128 patches per GPU,
loss = src/losses.py, gradient SUM all-reduce over the ranks (NCCL), fused unscale + clip(1.0) + AdamW(5e-4, wd 0).
This is NOT the config-4 number (tools/bench_train.py measures the whole hybrid step); it isolates the cost of the first
stage: 3 x 146.7 GFLOP x 32 = 14.1 TFLOP per step and GPU.

usage: python tools/bench_train_stage1.py [--batch 32] [--steps 5] [--warmup 2]
       python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/bench_train_stage1.py"""
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
from superresolution_b200.training import rrdbnet_forward  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=32)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=2)
    a = ap.parse_args()
    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', 0)))
    dev = torch.device('cuda', torch.cuda.current_device())
    if world > 1:
        dist.init_process_group('nccl')
    torch.manual_seed(0)
    model = HybridSuperResolutionModel(smoothing='none').to(dev)
    opt = FusedAdamW(model.stage1.parameters(), lr=5e-4, weight_decay=0.0)
    crit = CombinedLoss()
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    x = torch.rand(a.batch, 1, 128, 128, device=dev, generator=g)
    t = torch.rand(a.batch, 1, 256, 256, device=dev, generator=g) * 0.02
    t[torch.rand(t.shape, device=dev, generator=g) < 0.02] = 0.9  # ~2 % bright "stars": both loss branches fire

    ev = lambda: torch.cuda.Event(enable_timing=True)
    phases = {'forward+loss': 0.0, 'backward': 0.0, 'allreduce': 0.0, 'optimizer': 0.0}

    def step(record):
        e = [ev() for _ in range(5)]
        opt.zero_grad()
        e[0].record()
        loss, _ = crit(rrdbnet_forward(model.stage1, x, accumulate_into_grad=True), t)
        e[1].record()
        loss.backward()
        e[2].record()
        if world > 1:
            allreduce_gradients(opt.flat_grad)
        e[3].record()
        opt.step(grad_scale=1.0, max_norm=1.0, world=world)
        e[4].record()
        if record:
            torch.cuda.synchronize()
            for k, (i, j) in zip(phases, ((0, 1), (1, 2), (2, 3), (3, 4))):
                phases[k] += e[i].elapsed_time(e[j])
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
        flop = 3 * 146.71e9 * a.batch
        print(json.dumps({'workload': f'stage-1 (RRDBNet) training step, batch {a.batch}/GPU of 128x128, weighted-L1 loss, '
                                      f'grad all-reduce, clip 1.0, AdamW', 'n_gpus': world, 'ms_per_step': round(per, 2),
                          'samples_per_s': round(world * a.batch / (per / 1e3), 1),
                          'tflops_per_gpu': round(flop / (per / 1e3) / 1e12, 1),
                          'phase_ms': {k: round(v / a.steps, 2) for k, v in phases.items()},
                          'loss': round(loss.item(), 5),
                          'peak_mem_gb': round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
