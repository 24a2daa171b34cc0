"""Secondary workloads of bench.py (BASELINE.json configs[0], [2], [3], [4]): each returns a dict that
bench.py puts under ``extra`` of its single JSON line, with its own clock sample.  All GPU times are CUDA events on the
launching stream after warm-up, max over ranks; inputs are larger than L2 or L2 is flushed between iterations.

  config3   4096x4096 u16 frame -> 16384^2 u16, 1849 tiles of 128 px on stride 96, the tiles dealt out by count over the
            ranks of this run (tiling.infer_frame, seams='tiles': p2p exchange of partial sums, no collective)
  config4   hybrid training step, batch 32 per GPU, weighted-L1 loss, NCCL gradient all-reduce, clip 1.0, AdamW 5e-4
  config5   LOG normalise of 10 k HR+LR pairs, PSNR/SSIM of 10 k 512^2 pairs, robust statistics (rank 0 only)
  cpu       the fp32 oracle on the host cores at the reference's own size: config 1 (RRDBNet only, one 128^2 patch,
            median of 5) and one full hybrid 128^2 patch
"""
import json
import os
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
FLOP_PER_PATCH = 725.74e9


def _peaks():
    try:
        d = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        return float(d['hbm_gbs']), float(d['bf16_tflops']), float(d['bf16_tflops_sustained']), 'MEASURED_PEAKS.json'
    except Exception:
        return 6650.0, 1590.0, 1400.0, 'fallback (B200_PROFILING.md)'


def _ev():
    return torch.cuda.Event(enable_timing=True)


def _max_over_ranks(ms, dev, world):
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()
    return ms


def _timed(fn, warmup, iters):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = _ev(), _ev()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def config3(model, dev, rank, world, sampler_cls, gpu_index, size=4096):
    """Whole-field tiled inference sharded over the ranks of this run (strong scaling: one frame, N GPUs)."""
    import torch.distributed as dist
    from superresolution_b200 import tiling
    g = torch.Generator().manual_seed(2)
    frame = torch.randint(0, 65536, (size, size), generator=g, dtype=torch.int32).to(torch.uint16).to(dev)
    seams = 'tiles' if world > 1 else 'single'
    plan = tiling.tile_plan(size, size, 128, 96, 4, world)
    out = {}

    def run():
        out['y'], out['rows'] = tiling.infer_frame(model, frame, rank=rank, world=world)

    run()  # warm-up: workspace, NCCL p2p channels
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = sampler_cls(gpu_index)
    sampler.start()
    ts = []
    for _ in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = _ev(), _ev()
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ts.append(_max_over_ranks(e0.elapsed_time(e1), dev, world))
    clocks = sampler.stop()
    ms = min(ts)
    ntiles = sum(len(p['tiles']) for p in plan)
    del frame
    return {'workload': f'{size}x{size} u16 frame -> {4 * size}x{4 * size} u16, 128-px tiles on stride 96, halo blend',
            'scaling': 'strong', 'n_gpus': world, 'seams': seams, 'tiles': ntiles,
            'tiles_per_rank': [len(p['tiles']) for p in plan], 'frame_ms': round(ms, 2),
            'output_mpix_per_s': round(16 * size * size / 1e6 / (ms / 1e3), 2),
            'tiles_per_s': round(ntiles / (ms / 1e3), 1),
            'model_tflops': round(ntiles * FLOP_PER_PATCH / (ms / 1e3) / 1e12, 1),
            'timing': 'CUDA events around infer_frame (tile gather, forwards, blend, p2p exchange, finalise); best of 2 '
                      'after 1 warm-up; max over ranks', 'clocks': clocks}


def config4(dev, rank, world, sampler_cls, gpu_index, batch=32, micro=16, steps=3):
    """Hybrid training step (scripts/Modello_supporto.py:125-141 on libsrb) with the gradient all-reduce over the ranks."""
    import torch.distributed as dist
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.losses import CombinedLoss
    from superresolution_b200.optim import FusedAdamW, allreduce_gradients_start, allreduce_gradients_wait
    from superresolution_b200.training import _head, _tail, stage2_offset
    from superresolution_b200 import ops
    torch.manual_seed(0)
    model = HybridSuperResolutionModel(smoothing='none').to(dev).train()
    opt = FusedAdamW(model.parameters(), lr=5e-4, weight_decay=0.0)
    crit = CombinedLoss()
    cut = stage2_offset(model, opt)
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    x = torch.rand(batch, 1, 128, 128, device=dev, generator=g)
    t = torch.rand(batch, 1, 512, 512, device=dev, generator=g) * 0.02
    t[torch.rand(t.shape, device=dev, generator=g) < 0.02] = 0.9  # ~2 % bright "stars": both loss branches fire
    ar_evs = []

    def step(record):
        opt.zero_grad()
        y1 = _head(model, x, True)
        g1 = torch.empty_like(y1)
        loss = None
        with ops.weight_pack_cache():  # as train_step: one pack per weight and step
            for s in range(0, batch, micro):
                ys = y1[s:s + micro].detach().requires_grad_(True)
                loss, _ = crit(_tail(model, ys, 0.1), t[s:s + micro])
                (loss * (ys.shape[0] / batch)).backward()
                g1[s:s + micro] = ys.grad
        works = allreduce_gradients_start(opt.flat_grad, lo=cut) if world > 1 else []  # stage 2 is final: reduce it
        y1.backward(g1)                                                                # under the stage-1 backward
        e0, e1 = _ev(), _ev()
        e0.record()
        if world > 1:
            works += allreduce_gradients_start(opt.flat_grad, hi=cut)
            allreduce_gradients_wait(works)
        e1.record()
        opt.step(grad_scale=1.0, max_norm=1.0, world=world)
        if record:
            ar_evs.append((e0, e1))
        return loss

    step(False)  # two untimed steps: the second one runs with the caching allocator already grown (at 8 ranks the
    step(False)  # first timed step otherwise waits in the all-reduce for the rank that is still allocating)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = sampler_cls(gpu_index)
    sampler.start()
    e0, e1 = _ev(), _ev()
    e0.record()
    for _ in range(steps):
        loss = step(True)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    per = _max_over_ranks(e0.elapsed_time(e1), dev, world) / steps
    res = {'workload': f'hybrid (RRDBNet + HAT) training step, batch {batch}/GPU of 128x128, stage-2 micro-batches of '
                       f'{micro}, weighted-L1 loss, gradient all-reduce, clip 1.0, AdamW', 'scaling': 'weak',
           'n_gpus': world, 'ms_per_step': round(per, 1), 'samples_per_s': round(world * batch / (per / 1e3), 2),
           'model_tflops_per_gpu': round(3 * FLOP_PER_PATCH * batch / (per / 1e3) / 1e12, 1),
           'exposed_allreduce_ms': round(sum(a.elapsed_time(b) for a, b in ar_evs) / steps, 2), 'allreduce_bytes': opt.numel * 4,
           'loss': round(float(loss.item()), 5), 'peak_mem_gb': round(torch.cuda.max_memory_allocated() / 2 ** 30, 1),
           'clocks': clocks}
    del model, opt, x, t
    torch.cuda.empty_cache()
    return res


def config5(dev, sampler_cls, gpu_index, pairs=10000):
    """HBM-bound neighbours: achieved GB/s = algorithmic bytes / CUDA-event time, vs the measured copy peak."""
    from superresolution_b200.losses import CombinedLoss
    from superresolution_b200.metrics import _psnr_ssim
    from superresolution_b200.normalization import log_normalize_u16, robust_stats
    peak, _, _, src = _peaks()
    chunk = 500  # 500 x 1 MiB fp32 HR images = 524 MB per chunk: larger than the 126 MB L2
    nchunks = (pairs + chunk - 1) // chunk
    g = torch.Generator(device=dev).manual_seed(3)
    hr = torch.expm1(torch.randn(chunk, 512, 512, device=dev, generator=g) * 1.5 + 2.0)
    hr[torch.rand(hr.shape, device=dev, generator=g) < 0.001] = float('nan')
    hr[torch.rand(hr.shape, device=dev, generator=g) < 0.01] *= -1
    lr = torch.expm1(torch.randn(chunk, 128, 128, device=dev, generator=g) * 1.5 + 2.0)
    sampler = sampler_cls(gpu_index)
    sampler.start()
    res = {}

    def roof(ms, nbytes):
        gbs = nbytes / 1e9 / (ms / 1e3)
        return {'ms': round(ms, 3), 'roofline': {'bound': 'hbm', 'achieved': round(gbs, 1), 'peak': peak, 'unit': 'GB/s',
                                                  'frac': round(gbs / peak, 3), 'peak_source': src}}

    def norm():
        for _ in range(nchunks):
            log_normalize_u16(hr, 0.5, 9.0)
            log_normalize_u16(lr, 0.5, 9.0)
    ms = _timed(norm, 1, 3)
    res['log_normalize'] = dict(roof(ms, nchunks * (hr.numel() + lr.numel()) * 6), pairs=nchunks * chunk, bytes_per_px=6)

    def rstats():
        robust_stats([hr] * nchunks, 3.0)
    ms = _timed(rstats, 1, 2)
    res['robust_stats'] = dict(roof(ms, nchunks * hr.numel() * 12), images=nchunks * chunk, bytes_per_px=12,
                               note='three 4-byte passes (exact 3-level radix select), 3 host syncs inside')
    del hr, lr
    t = torch.rand(chunk, 1, 512, 512, device=dev, generator=g)
    p = (t + 0.01 * torch.randn(t.shape, device=dev, generator=g)).clamp_(0, 1)

    def metrics():
        for _ in range(nchunks):
            for i in range(0, chunk, 250):
                _psnr_ssim(p[i:i + 250], t[i:i + 250])
    ms = _timed(metrics, 1, 3)
    res['psnr_ssim'] = dict(roof(ms, nchunks * t.numel() * 8), pairs=nchunks * chunk, bytes_per_px=8)
    # weighted-L1 loss + gradient kernel (srb_weighted_l1 through the C ABI, caller-owned buffers): 250 pairs = 786 MB of
    # traffic per call, larger than L2; and the same loss through the autograd module (torch allocations + the
    # grad_output scaling kernel inside the timed region) on the training batch shape
    from superresolution_b200 import _lib as L
    lib = L.lib()
    pk, tk = p[:250].contiguous(), t[:250].contiguous()
    gk = torch.empty_like(pk)
    lossbuf = torch.zeros(1, device=dev)
    scr = torch.empty(lib.srb_loss_scratch_bytes(), dtype=torch.uint8, device=dev)

    def l1_kernel():
        L.check(lib.srb_weighted_l1(L.ptr(pk), L.ptr(tk), pk.numel(), L.ptr(lossbuf), L.ptr(gk), L.ptr(scr),
                                    L.current_stream()), 'weighted_l1')
    ms = _timed(l1_kernel, 2, 7)
    res['weighted_l1_kernel'] = dict(roof(ms, pk.numel() * 12), pairs=250, bytes_per_px=12,
                                     note='loss + dLoss/dpred in one pass: read pred, target; write grad')
    del pk, tk, gk
    crit = CombinedLoss()
    pred = p[:32].clone().requires_grad_(True)
    big = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def loss():
        big.zero_()  # L2 flush: the 64 MB working set would otherwise stay resident
        pred.grad = None
        l, _ = crit(pred, t[:32])
        l.backward()
    ms = _timed(loss, 2, 7) - _timed(big.zero_, 2, 7)
    res['weighted_l1_fwd_bwd'] = dict(roof(ms, pred.numel() * 12), shape=[32, 1, 512, 512], bytes_per_px=12,
                                      note='autograd module path (allocations, grad_output scaling kernel); L2 flushed, '
                                           'flush time subtracted')
    res['clocks'] = sampler.stop()
    del t, p, big
    torch.cuda.empty_cache()
    return res


def cpu_baselines():
    """The fp32 oracle port of the reference on the box's host cores at the reference's own patch size (128x128):
    config 1 = RRDBNet only (BASELINE.json configs[0]), and one full hybrid patch (configs[1] per patch)."""
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    o = Oracle(output_size=512, smoothing='balanced').eval()
    g = torch.Generator().manual_seed(1)
    x = torch.randint(0, 65536, (1, 1, 128, 128), generator=g).float() / 65535.0
    with torch.no_grad():
        o.stage1(x)
        ts = []
        for _ in range(5):
            t0 = time.perf_counter()
            o.stage1(x)
            ts.append(time.perf_counter() - t0)
        ts.sort()
        s1 = ts[2]
        t0 = time.perf_counter()
        o(x)
        full = time.perf_counter() - t0
    return {'cores': cores, 'kind': 'port',
            'config1_rrdbnet_fp32_128px': {'s_per_patch': round(s1, 3), 'output_mpix_per_s': round(0.065536 / s1, 4),
                                           'gflops': round(146.71 / s1, 1), 'sample': 'median of 5 after 1 warm-up'},
            'full_hybrid_fp32_128px': {'s_per_patch': round(full, 3), 'output_mpix_per_s': round(0.262144 / full, 4),
                                       'gflops': round(725.74 / full, 1), 'sample': 'one 128x128 patch after the stage-1 warm-up'}}
