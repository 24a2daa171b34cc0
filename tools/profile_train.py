"""Kernel-time table of one training micro-batch (hybrid forward + backward) from torch.profiler (CUPTI):
which libsrb / ATen kernels the config-4 step spends its time in.  usage: python tools/profile_train.py [--micro 4]"""
import argparse
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel  # noqa: E402
from superresolution_b200.losses import CombinedLoss  # noqa: E402
from superresolution_b200.optim import FusedAdamW  # noqa: E402
from superresolution_b200.training import hybrid_forward  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--micro', type=int, default=4)
    ap.add_argument('--rows', type=int, default=45)
    a = ap.parse_args()
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    model = HybridSuperResolutionModel(smoothing='none').to(dev).train()
    opt = FusedAdamW(model.parameters(), lr=5e-4, weight_decay=0.0)
    crit = CombinedLoss()
    x = torch.rand(a.micro, 1, 128, 128, device=dev)
    t = torch.rand(a.micro, 1, 512, 512, device=dev) * 0.02

    def one():
        opt.zero_grad()
        loss, _ = crit(hybrid_forward(model, x, accumulate_into_grad=True), t)
        loss.backward()

    one()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        one()
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=a.rows, max_name_column_width=90))


if __name__ == '__main__':
    main()
