"""conv_tc phase probe (library built with `make EXTRA=-DSRB_CONV_PHASES`): one conv at the bench geometry.
usage: python tools/conv_probe.py Cin Cout [taps=9] [act=0|2] [R=256] [B=16] [mode=plain|pre|alpha|proj|gelugrad]"""
import os, sys, ctypes as C, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200 import _lib as L, ops
cin, cout = int(sys.argv[1]), int(sys.argv[2])
taps = int(sys.argv[3]) if len(sys.argv) > 3 else 9
act = int(sys.argv[4]) if len(sys.argv) > 4 else 0
R = int(sys.argv[5]) if len(sys.argv) > 5 else 256
B = int(sys.argv[6]) if len(sys.argv) > 6 else 16
mode = sys.argv[7] if len(sys.argv) > 7 else 'plain'
torch.manual_seed(0)
k = 3 if taps == 9 else 1
x = torch.randn(B, R, R, cin, device='cuda').bfloat16()
w = ops.pack_conv_weight(torch.randn(cout, cin, k, k, device='cuda') * 0.05)
bias = torch.zeros(cout, device='cuda')
out = torch.empty(B, R, R, cout, device='cuda', dtype=torch.bfloat16)
lib = L.lib()
mkb = lambda c: torch.randn(B, R, R, c, device='cuda').bfloat16()
if mode != 'plain':  # training epilogues (conv_nhwc_train)
    kw = {'pre': dict(pre_out=mkb(cout)), 'alpha': dict(res1=mkb(cout), alpha1_samples=torch.ones(B, device='cuda')),
          'proj': dict(res1=mkb(cout), alpha1_samples=torch.ones(B, device='cuda'), res2=mkb(cout),
                       res2_cs=torch.full((B, cout), 0.01, device='cuda')),
          'gelugrad': dict(gelu_grad_of=mkb(cout))}[mode]
    _plain = ops.conv_nhwc
    ops.conv_nhwc = lambda x, cin, w, bias, out=None, act=0: ops.conv_nhwc_train(x, cin, w, bias, out=out, act=act, **kw)
lib.srb_debug_set_phase_buffer.argtypes = [C.c_void_p]; lib.srb_debug_set_phase_buffer.restype = None
buf = torch.zeros(148 * 16, dtype=torch.int64, device='cuda')
for _ in range(3):
    ops.conv_nhwc(x, cin, w, bias, out=out, act=act)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.conv_nhwc(x, cin, w, bias, out=out, act=act)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 100
lib.srb_debug_set_phase_buffer(C.c_void_p(buf.data_ptr()))
ops.conv_nhwc(x, cin, w, bias, out=out, act=act)
torch.cuda.synchronize()
lib.srb_debug_set_phase_buffer(None)
p = buf.view(148, 16).double().mean(0)
fl = 2.0 * B * R * R * cin * cout * taps
print(f'conv {cin}->{cout} taps {taps} act {act} B {B} R {R}: {us:.1f} us/launch (warm L2, back to back), {fl / us / 1e6:.0f} TFLOP/s')
names = ['mma_total', 'wait_acc_empty', 'wait_patch', 'wait_weight', 'epi_total', 'epi_wait_acc_full', 'epi_tmem_ld',
         'epi_math', 'epi_bar1(+read wait)', 'epi_pack+sts+fence', 'epi_bar2']
for n, v in zip(names, p[:11]):
    print(f'  {n:18s} {v.item() / 1e3:9.1f} kcyc')
print(f'  mma issue+other    {(p[0] - p[1] - p[2] - p[3]).item() / 1e3:9.1f} kcyc;  epilogue work {(p[4] - p[5]).item() / 1e3:9.1f} kcyc')
