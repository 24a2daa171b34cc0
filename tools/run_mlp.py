"""Run the fused MLP branch alone at the bench geometry (1M tokens) — target for `ncu --set full -k regex:mlp_tc`."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200 import ops
ntok, B = 16 * 256 * 256, 16
dev = 'cuda'
g = torch.Generator().manual_seed(0)
t = torch.randn(ntok, 96, device=dev).to(torch.bfloat16)
cab = torch.randn(ntok, 96, device=dev).to(torch.bfloat16)
cs = torch.rand(B, 96, device=dev) * 0.02
ln_g, ln_b = torch.ones(96, device=dev), torch.zeros(96, device=dev)
b1, b2 = torch.zeros(192, device=dev), torch.zeros(96, device=dev)
w1, b1 = ops.fold_ln_linear(torch.randn(192, 96, device=dev) / 10, b1, ln_g, ln_b)
w2 = ops.pack_conv_weight(torch.randn(96, 192, device=dev) / 14)
out = torch.empty_like(t)
for _ in range(3):
    ops.mlp_branch(t, w1, b1, w2, b2, cab=cab, cs=cs, out=out, tok_per_img=ntok // B)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.mlp_branch(t, w1, b1, w2, b2, cab=cab, cs=cs, out=out, tok_per_img=ntok // B)
e1.record(); torch.cuda.synchronize()
print('mlp_branch us/launch', e0.elapsed_time(e1) * 100)

import ctypes as C
from superresolution_b200 import _lib as L
lib = L.lib()
buf = torch.zeros(148 * 8, dtype=torch.int64, device='cuda')
lib.srb_debug_set_phase_buffer.argtypes = [C.c_void_p]; lib.srb_debug_set_phase_buffer.restype = None
lib.srb_debug_set_phase_buffer(C.c_void_p(buf.data_ptr()))
ops.mlp_branch(t, w1, b1, w2, b2, cab=cab, cs=cs, out=out, tok_per_img=ntok // B)
torch.cuda.synchronize()
lib.srb_debug_set_phase_buffer(None)
p = buf.view(148, 8).double().mean(0)
names = ['wait_d1', 'gelu+bar', 'ln_combine', 'ln_stats+A1', 'ln_fence', 'issue_loads', 'wait_d2+epi', 'epi_bar+copyout']
tot = p.sum().item() or 1.0  # all zero unless libsrb was built with EXTRA=-DSRB_MLP_PHASES
for n, v in zip(names, p):
    print(f'{n:12s} {v.item() / 55.35:8.0f} cyc/tile {100 * v.item() / tot:5.1f}%')
print('total cyc/tile', tot / 55.35)
