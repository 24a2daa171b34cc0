"""Secondary measurements for BASELINE.json configs 3 and 5 and the loss kernel.  `bench.py` keeps the
driver contract (config 2); this prints one JSON object per workload, all timed with CUDA events after warm-up:

  frame   config 3: synthetic 4096x4096 u16 frame, 128-px tiles on stride 96 (1849 tiles), 4x output 16384^2 u16,
          through superresolution_b200.tiling.infer_frame on this rank's band (torchrun -> one band per rank).
  aux     config 5: LOG normalise of HR [512,512] + LR [128,128] fp32 pairs and PSNR/SSIM of [1,1,512,512] pairs,
          10 000 pairs in chunks larger than L2; weighted-L1 loss forward+gradient on [32,1,512,512].
          HBM-bound kernels: achieved GB/s = algorithmic bytes (6 B/px LOG-norm, 8 B/px PSNR+SSIM,
          8 B read + 4 B grad write per px loss) / time, against MEASURED_PEAKS.json hbm_gbs.

usage: python tools/bench_configs.py [frame] [aux] [--frame-size 4096] [--pairs 10000]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200 import tiling  # noqa: E402
from superresolution_b200.architecture import HybridSuperResolutionModel  # noqa: E402
from superresolution_b200.losses import CombinedLoss  # noqa: E402
from superresolution_b200.metrics import _psnr_ssim  # noqa: E402
from superresolution_b200.normalization import log_normalize_u16, robust_stats  # noqa: E402
from superresolution_b200.optim import FusedAdamW  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']), 'MEASURED_PEAKS.json'
    except Exception:
        return 6454.0, 'fallback (B200_PROFILING.md)'


def timed(fn, warmup=2, iters=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def bench_frame(size, rank, world, seams='recompute'):
    dev = torch.device('cuda', torch.cuda.current_device())
    torch.manual_seed(0)
    model = HybridSuperResolutionModel(output_size=512).to(dev).eval()
    g = torch.Generator().manual_seed(2)
    frame = torch.randint(0, 65536, (size, size), generator=g, dtype=torch.int32).to(torch.uint16).to(dev)
    ntiles = len(tiling.plan(size, size))
    out = {}

    def run():
        out['y'], out['band'] = tiling.infer_frame(model, frame, rank=rank, world=world, seams=seams)

    ms = timed(run, warmup=1, iters=3)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = t.item()
    r0, r1 = out['band']
    if rank == 0:
        print(json.dumps({'workload': f'config3: {size}x{size} u16 frame, 128-px tiles stride 96, 4x output u16',
                          'n_gpus': world, 'seams': seams, 'tiles': ntiles, 'frame_ms': round(ms, 2),
                          'output_mpix_per_s': round(16 * size * size / 1e6 / (ms / 1e3), 2),
                          'tiles_per_s': round(ntiles / (ms / 1e3), 1), 'band_rows_rank0': [r0, r1],
                          'timing': 'CUDA events around infer_frame (tile gather, forwards, blend, finalise); '
                                    'median of 3 after 1 warm-up; max over ranks'}))


def bench_aux(pairs):
    dev = torch.device('cuda', torch.cuda.current_device())
    peak, src = hbm_peak()
    chunk = 500  # 500 x 1 MiB fp32 HR images = 524 MB per chunk: larger than the 126 MB L2
    g = torch.Generator(device=dev).manual_seed(3)
    # heavy-tailed positive counts with 0.1 % NaN and 1 % negatives
    hr = torch.expm1(torch.randn(chunk, 512, 512, device=dev, generator=g) * 1.5 + 2.0)
    hr[torch.rand(hr.shape, device=dev, generator=g) < 0.001] = float('nan')
    hr[torch.rand(hr.shape, device=dev, generator=g) < 0.01] *= -1
    lr = torch.expm1(torch.randn(chunk, 128, 128, device=dev, generator=g) * 1.5 + 2.0)
    nchunks = (pairs + chunk - 1) // chunk

    def norm():
        for _ in range(nchunks):
            log_normalize_u16(hr, 0.5, 9.0)
            log_normalize_u16(lr, 0.5, 9.0)

    ms = timed(norm)
    px = nchunks * (hr.numel() + lr.numel())
    gbs = px * 6 / 1e9 / (ms / 1e3)
    print(json.dumps({'workload': f'config5a: LOG normalise {nchunks * chunk} HR[512,512]+LR[128,128] fp32 pairs -> u16',
                      'ms': round(ms, 3), 'pairs_per_s': round(nchunks * chunk / (ms / 1e3)),
                      'roofline': {'bound': 'hbm', 'achieved': round(gbs, 1), 'peak': peak, 'unit': 'GB/s',
                                   'frac': round(gbs / peak, 3), 'algorithmic_bytes_per_px': 6, 'peak_source': src}}))

    # the step before the normalisation: exact black / white percentiles of the HR population
    stats = {}

    def rstats():
        stats['v'] = robust_stats([hr] * nchunks, 3.0)

    ms = timed(rstats, warmup=1, iters=3)
    px = nchunks * hr.numel()
    gbs = px * 12 / 1e9 / (ms / 1e3)
    print(json.dumps({'workload': f'config5c: robust stats (3rd / 99.99th percentile of log1p over the valid pixels) of '
                                  f'{nchunks * chunk} HR[512,512] fp32 images, exact 3-level radix select',
                      'ms': round(ms, 3), 'gmin_gmax': [float(stats['v'][0]), float(stats['v'][1])],
                      'roofline': {'bound': 'hbm', 'achieved': round(gbs, 1), 'peak': peak, 'unit': 'GB/s',
                                   'frac': round(gbs / peak, 3), 'algorithmic_bytes_per_px': 12,
                                   'note': 'three 4-byte reads per pixel (one per radix level); includes 3 host syncs',
                                   'peak_source': src}}))
    del hr, lr

    t = torch.rand(chunk, 1, 512, 512, device=dev, generator=g)
    p = (t + 0.01 * torch.randn(t.shape, device=dev, generator=g)).clamp_(0, 1)
    B = 250  # pairs per Metrics.update call

    def metrics():
        for _ in range(nchunks):
            for i in range(0, chunk, B):
                _psnr_ssim(p[i:i + B], t[i:i + B])

    ms = timed(metrics)
    px = nchunks * t.numel()
    gbs = px * 8 / 1e9 / (ms / 1e3)
    print(json.dumps({'workload': f'config5b: PSNR+SSIM of {nchunks * chunk} [1,1,512,512] fp32 pairs ({B} per call)',
                      'ms': round(ms, 3), 'pairs_per_s': round(nchunks * chunk / (ms / 1e3)),
                      'roofline': {'bound': 'hbm', 'achieved': round(gbs, 1), 'peak': peak, 'unit': 'GB/s',
                                   'frac': round(gbs / peak, 3), 'algorithmic_bytes_per_px': 8, 'peak_source': src}}))

    crit = CombinedLoss()
    pred = p[:32].clone().requires_grad_(True)
    tgt = t[:32]
    big = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def loss():
        big.zero_()  # L2 flush: the 64 MB working set would otherwise stay resident
        pred.grad = None
        l, _ = crit(pred, tgt)
        l.backward()

    def flush_only():
        big.zero_()

    ms = timed(loss, iters=9) - timed(flush_only, iters=9)
    gbs = pred.numel() * 12 / 1e9 / (ms / 1e3)
    print(json.dumps({'workload': 'weighted-L1 loss forward + gradient, [32,1,512,512] fp32 (L2 flushed, flush time subtracted)',
                      'ms': round(ms, 4),
                      'roofline': {'bound': 'hbm', 'achieved': round(gbs, 1), 'peak': peak, 'unit': 'GB/s',
                                   'frac': round(gbs / peak, 3), 'algorithmic_bytes_per_px': 12, 'peak_source': src}}))


def bench_optim():
    """config 4's optimizer half: unscale + clip-norm + AdamW over the model's 20.8 M fp32 parameters (333 MB of state:
    larger than L2).  Algorithmic bytes: 4 (norm pass) + 16 read + 12 write = 32 B per parameter."""
    dev = torch.device('cuda', torch.cuda.current_device())
    peak, src = hbm_peak()
    torch.manual_seed(0)
    model = HybridSuperResolutionModel(output_size=512, smoothing='none').to(dev)
    opt = FusedAdamW(model.parameters(), lr=5e-4, weight_decay=0.0)
    opt.flat_grad.normal_()
    ms = timed(lambda: opt.step(grad_scale=65536.0, max_norm=1.0), warmup=3, iters=11)
    gbs = opt.numel * 32 / 1e9 / (ms / 1e3)
    print(json.dumps({'workload': f'config4 optimizer half: grad-norm + unscale + clip + AdamW, {opt.numel} fp32 parameters',
                      'ms': round(ms, 4),
                      'roofline': {'bound': 'hbm', 'achieved': round(gbs, 1), 'peak': peak, 'unit': 'GB/s',
                                   'frac': round(gbs / peak, 3), 'algorithmic_bytes_per_param': 32, 'peak_source': src}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('what', nargs='*', default=['frame', 'aux'])
    ap.add_argument('--frame-size', type=int, default=4096)
    ap.add_argument('--pairs', type=int, default=10000)
    ap.add_argument('--seams', default='recompute', choices=['recompute', 'p2p'])
    a = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', 0)))
    if world > 1:
        torch.distributed.init_process_group('nccl')
    if 'aux' in a.what and rank == 0:
        bench_aux(a.pairs)
        bench_optim()
    if 'frame' in a.what:
        bench_frame(a.frame_size, rank, world, a.seams)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == '__main__':
    main()
