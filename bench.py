#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json: "SR output Mpix/s at 1/2/4/8 B200 (bf16)").

Workload (configs[1]): full HybridSuperResolutionModel (RRDBNet 2x + HAT 2x), bf16 tensor-core math with fp32
accumulation, batch 16 of synthetic 128x128 single-channel 16-bit patches per GPU, smoothing='balanced',
random-init weights (seed 0).  One "step" = one forward of the batch.  Multi-GPU = independent tile batches per rank
(weak scaling, no data-path collective).

  python bench.py --gpus N --steps K --warmup W          (torchrun launches it for N>1)
  python bench.py --impl reference ...                   CPU arm: the oracle port of the reference on host cores

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'sr_output_mpix_per_s'
UNIT = 'Mpix/s'
BATCH, PATCH = 16, 128
FLOP_PER_PATCH = 725.74e9  # 362.869 GMAC per 128x128 patch
WORKLOAD = 'hybrid_rrdb23+hat96_bf16_infer_b16_128px'


def load_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get('bf16_tflops_sustained', 1400.0), d.get('hbm_gbs', 6650.0), 'measured (MEASURED_PEAKS.json, sustained)'
    return 1400.0, 6650.0, 'fallback (B200_PROFILING.md)'


def synthetic_batch(B, size, seed=1):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 65536, (B, 1, size, size), generator=g).float() / 65535.0


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonSwPowerCap: 'sw_power_cap',
                 nv.nvmlClocksThrottleReasonHwSlowdown: 'hw_slowdown',
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: 'sw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: 'hw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: 'hw_power_brake'}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {'sm_mhz': s[len(s) // 2] if s else None, 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons)}


def visible_gpu_index(local_rank):
    vis = os.environ.get('CUDA_VISIBLE_DEVICES')
    if vis:
        try:
            return int(vis.split(',')[local_rank])
        except Exception:
            return local_rank
    return local_rank


def cpu_oracle_throughput(steps, warmup, size=PATCH):
    """Reference arm / cpu_baseline: the oracle port of the reference model (fp32, eager, no_grad) on all host cores,
    on a bounded sample of the workload: ONE size x size patch per step instead of the batch of 16 (the reference's own
    patch size, 128; ~3 s per patch on 16 cores, per-patch cost is batch-independent in eager fp32)."""
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    o = Oracle(output_size=4 * size, smoothing='balanced').eval()
    x = synthetic_batch(1, size)
    with torch.no_grad():
        for _ in range(warmup):
            o(x)
        t0 = time.perf_counter()
        for _ in range(steps):
            o(x)
        dt = time.perf_counter() - t0
    mpix = steps * (4 * size) ** 2 / 1e6
    return mpix / dt, dt / steps * 1e3, cores, f'{steps} x one {size}x{size} patch (fp32 oracle port, torch eager)'


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    v, ms, cores, sample = cpu_oracle_throughput(args.steps, args.warmup)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': WORKLOAD, 'note': 'reference model cannot be imported (BasicSR/HAT not vendored): '
                                                 'oracle port on host cores'},
        'cpu_baseline': {'value': v, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': v, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line), flush=True)


def run_gpu(args):
    import torch.distributed as dist
    from superresolution_b200.architecture import HybridSuperResolutionModel

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device: libsrb has no CPU path')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    torch.manual_seed(0)
    model = HybridSuperResolutionModel(output_size=4 * PATCH, smoothing='balanced').to(dev).eval()
    x_host = synthetic_batch(BATCH, PATCH, seed=1 + rank).pin_memory()
    x_dev = x_host.to(dev)
    out_mpix = BATCH * (4 * PATCH) ** 2 / 1e6

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        model.pack_weights()
        y = model(x_dev)  # first call: workspace allocation
        torch.cuda.synchronize()
        # ---- capture the forward in a CUDA graph (stream-ordered, allocation-free inside libsrb) ----
        graph = None
        if not args.no_graph:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                model(x_dev)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                y_graph = model(x_dev)

        def step():
            if graph is not None:
                graph.replay()
                return y_graph
            return model(x_dev)

        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
        for _ in range(max(args.warmup, 3)):
            step()
        barrier()
        sampler = ClockSampler(visible_gpu_index(local_rank))
        sampler.start()
        # ---- value: device-resident inputs, one CUDA-event pair per step, L2 flushed between steps ----
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for e0, e1 in evs:
            flush.zero_()
            e0.record()
            step()
            e1.record()
        torch.cuda.synchronize()
        dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in evs)
        # ---- e2e: public API with host buffers, H2D + forward + D2H inside the timed region ----
        y_host = torch.empty(BATCH, 1, 4 * PATCH, 4 * PATCH, dtype=torch.float32).pin_memory()
        # the package's host-to-host entry point: every step's H2D copy, forward and D2H copy are issued inside the
        # timed region; the copies of neighbouring steps run on their own streams under the forward
        from superresolution_b200.pipeline import infer_host_batches
        infer_host_batches(model, [x_host] * 2, out=[y_host] * 2)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        infer_host_batches(model, [x_host] * args.steps, out=[y_host] * args.steps)
        e1.record()
        torch.cuda.synchronize()
        e2e_ms = e0.elapsed_time(e1)
        clocks = sampler.stop()
        # ---- per-kernel-family CUDA-event times (every launch bracketed on the launching stream).  A ~10 ms device
        # sleep is queued first so that the host has enqueued the launches before the GPU reaches them: the event
        # deltas then hold GPU time only, not host launch gaps.
        torch.cuda._sleep(20_000_000)
        prof = model.profile(x_dev)
        launches_per_step = model.launch_count(BATCH, PATCH, PATCH)
        graphed = graph is not None
        del flush, graph

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = t.tolist()
    # ---- secondary workloads (own clock samples): the sharded frame and the data-parallel train step run on every rank
    extra = {}
    if not args.no_extra:
        import bench_extra as X
        gi = visible_gpu_index(local_rank)
        extra['config3'] = X.config3(model, dev, rank, world, ClockSampler, gi)
        del model
        torch.cuda.empty_cache()
        extra['config4'] = X.config4(dev, rank, world, ClockSampler, gi)
        if rank == 0 and world == 1:
            extra['config5'] = X.config5(dev, ClockSampler, gi)
    if rank == 0:
        line = report(args, world, dev_ms, e2e_ms, prof, clocks, launches_per_step, x_host, y_host, graphed)
        if extra:
            line['extra'] = extra
        if world == 1 and not args.no_cpu_baseline:
            import bench_extra as X
            cb = X.cpu_baselines()
            full = cb['full_hybrid_fp32_128px']
            line['cpu_baseline'] = {'value': full['output_mpix_per_s'], 'unit': UNIT, 'cores': cb['cores'], 'kind': 'port',
                                    'sample': 'one full hybrid 128x128 patch (fp32 oracle port, torch eager, all host cores) '
                                              'instead of the batch of 16', 'config1_rrdbnet_fp32_128px':
                                        cb['config1_rrdbnet_fp32_128px'], 'full_hybrid_fp32_128px': full}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# name of the kernel behind each tensor-core profile class of srb_forward_profile, algorithmic FLOPs counted per launch
KERNELS = {'conv3x3_tcgen05': 'conv_tc_kernel<*,*,9> (tcgen05 implicit-GEMM 3x3 conv)',
           'rrdb_chain_tcgen05': 'conv_chain_kernel (one RRDB = 15 dense-block convs per launch, tcgen05)',
           'linear_tcgen05': 'conv_tc_kernel<*,*,1> (tcgen05 linear)',
           'hab_attention_tcgen05': 'hab_tc_kernel (LN1 + qkv + (S)W-MSA + proj; every contraction on tcgen05 / TMEM)',
           'mlp_tcgen05': 'mlp_tc_kernel (LN2 + fc1 + GELU + fc2 + residual; tcgen05, TMEM)',
           'ocab_attention': 'ocab_attn_kernel (mma.sync)'}


def report(args, world, dev_ms, e2e_ms, prof, clocks, launches, x_host, y_host, graphed):
    peak_sus, peak_hbm, _ = load_peaks()
    try:
        peak_burst = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['bf16_tflops'])
        peak_src = 'measured (MEASURED_PEAKS.json)'
    except Exception:
        peak_burst, peak_src = 1590.0, 'fallback (B200_PROFILING.md)'
    # the clock record decides the denominator: no power cap seen -> the burst figure, else the sustained one
    capped = 'sw_power_cap' in (clocks.get('reasons') or [])
    peak_tf = peak_sus if capped else peak_burst
    out_mpix = BATCH * (4 * PATCH) ** 2 / 1e6
    total_prof_ms = sum(v[0] for v in prof.values())
    dom = max(KERNELS, key=lambda k: prof[k][0])  # dominant KERNEL of the step (largest sum of CUDA-event time)
    d_ms, d_flops, d_n = prof[dom]
    achieved = d_flops / (d_ms * 1e-3) / 1e12 if d_ms > 0 else 0.0
    tc_ms = sum(prof[k][0] for k in KERNELS)
    tc_flops = sum(prof[k][1] for k in KERNELS)
    traffic, traffic_src = family_traffic(dom)
    return {
        'metric': METRIC, 'value': world * out_mpix * args.steps / (dev_ms * 1e-3), 'unit': UNIT,
        'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': dev_ms / args.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
        'config': {'workload': WORKLOAD, 'batch_per_gpu': BATCH, 'patch': PATCH, 'smoothing': 'balanced',
                   'weights': 'random-init seed 0', 'cuda_graph': graphed,
                   'l2': 'flushed between timed steps (256 MiB write)', 'patches_per_s':
                       world * BATCH * args.steps / (dev_ms * 1e-3),
                   'model_tflops': world * BATCH * FLOP_PER_PATCH * args.steps / (dev_ms * 1e-3) / 1e12},
        'clocks': clocks,
        'e2e': {'value': world * out_mpix * args.steps / (e2e_ms * 1e-3), 'unit': UNIT,
                'h2d_bytes_per_step': x_host.numel() * 4, 'd2h_bytes_per_step': y_host.numel() * 4},
        'gpu_launches': launches * args.steps,
        'roofline': {'bound': 'tensor', 'achieved': achieved, 'peak': peak_tf, 'unit': 'TFLOP/s',
                     'frac': achieved / peak_tf, 'frac_of_burst': achieved / peak_burst,
                     'frac_of_sustained': achieved / peak_sus, 'traffic': traffic, 'traffic_source': traffic_src,
                     'peak_source': peak_src + (': sustained (sw_power_cap seen)' if capped else ': burst (no power cap in '
                                                                                              'the clock record)'),
                     'kernel': KERNELS[dom], 'launches_per_step': d_n, 'kernel_ms_per_step': d_ms,
                     'share_of_step': d_ms / total_prof_ms if total_prof_ms else None,
                     'whole_step': {'achieved': BATCH * FLOP_PER_PATCH / (dev_ms / args.steps * 1e-3) / 1e12,
                                    'frac': BATCH * FLOP_PER_PATCH / (dev_ms / args.steps * 1e-3) / 1e12 / peak_tf},
                     'all_tensor_kernels': {'achieved': tc_flops / (tc_ms * 1e-3) / 1e12 if tc_ms else 0.0,
                                            'frac': tc_flops / (tc_ms * 1e-3) / 1e12 / peak_tf if tc_ms else 0.0,
                                            'ms_per_step': tc_ms},
                     'by_kernel_ms': {k: round(v[0], 4) for k, v in prof.items()},
                     'by_kernel_tflops': {k: round(v[1] / (v[0] * 1e-3) / 1e12, 1) for k, v in prof.items()
                                          if v[0] > 0 and v[1] > 0},
                     'by_kernel_frac': {k: round(v[1] / (v[0] * 1e-3) / 1e12 / peak_tf, 3) for k, v in prof.items()
                                        if v[0] > 0 and v[1] > 0}},
    }


# kernel-name substrings of each tensor-core family in the committed ncu capture (profiles/r02_kernel_traffic.json)
FAMILY_KERNELS = {'conv3x3_tcgen05': (', 9, 0',), 'rrdb_chain_tcgen05': ('conv_chain_kernel',), 'linear_tcgen05': (', 1, 0',),
                  'hab_attention_tcgen05': ('hab_tc_kernel',), 'mlp_tcgen05': ('mlp_tc_kernel',),
                  'ocab_attention': ('ocab_attn_kernel',)}


def family_traffic(family):
    """DRAM bytes (read + write) per launch of the family's kernels, averaged over its launches in one forward, from
    the committed ncu capture of tools/run_forward_once.py (dram__bytes_read.sum + dram__bytes_write.sum per launch;
    tools/ncu_metrics_table.py --json).  None when the file is absent."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'r02_kernel_traffic.json')
    try:
        tab = json.load(open(path))
    except Exception:
        return None, None
    tot = n = 0
    for name, row in tab.items():
        if any(k in name for k in FAMILY_KERNELS.get(family, ())):
            tot += row['dram_bytes_per_launch'] * row['launches']
            n += row['launches']
    return (tot / n if n else None), 'profiles/r02_kernel_traffic.json (ncu dram__bytes_read+write per launch, one forward)'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='srb', choices=['srb', 'reference'])
    ap.add_argument('--no-graph', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extra', action='store_true', help='skip extra.config3 / config4 / config5 (config 2 only)')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == '__main__':
    main()
