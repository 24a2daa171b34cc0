"""LayerNorm forward / backward and channel-attention backward alone at the training geometry (262144 tokens x 96), warm
and with the L2 flushed between launches — the target of `ncu --set full -k regex:layernorm96`."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200 import ops

ntok = 16 * 256 * 256
dev = 'cuda'
torch.manual_seed(0)
x = torch.randn(ntok, 96, device=dev).to(torch.bfloat16)
g = torch.randn(ntok, 96, device=dev).to(torch.bfloat16)
res = torch.randn(ntok, 96, device=dev).to(torch.bfloat16)
gamma, beta = torch.rand(96, device=dev) + 0.5, torch.randn(96, device=dev) * 0.1
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(name, fn, nbytes, cold):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    n = 10
    for _ in range(n):
        if cold:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    us = tot / n * 1e3
    print(f'{name:28s} {"cold" if cold else "warm"}  {us:8.1f} us  {nbytes / us / 1e3:7.0f} GB/s')


for cold in (False, True):
    timed('layernorm96', lambda: ops.layernorm96(x, gamma, beta), ntok * 96 * 4, cold)
    timed('layernorm96_bwd(res bf16)', lambda: ops.layernorm96_bwd(x, g, gamma, res), ntok * 96 * 8, cold)
    timed('layernorm96_bwd(no res)', lambda: ops.layernorm96_bwd(x, g, gamma), ntok * 96 * 6, cold)
