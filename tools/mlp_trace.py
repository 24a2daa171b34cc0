"""Timeline of mlp_tc_kernel (build with EXTRA=-DSRB_MLP_TRACE): clock64 stamps of CTA 0, its tiles 3 and 4, per warp."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200 import ops, _lib as L
ntok, B = 16 * 256 * 256, 16
dev = 'cuda'
g = torch.Generator().manual_seed(0)
t = torch.randn(ntok, 96, device=dev).to(torch.bfloat16)
cab = torch.randn(ntok, 96, device=dev).to(torch.bfloat16)
cs = torch.rand(B, 96, device=dev) * 0.02
w1, b1 = ops.fold_ln_linear(torch.randn(192, 96, device=dev) / 10, torch.zeros(192, device=dev), torch.ones(96, device=dev), torch.zeros(96, device=dev))
w2 = ops.pack_conv_weight(torch.randn(96, 192, device=dev) / 14)
b2 = torch.zeros(96, device=dev)
out = torch.empty_like(t)
lib = L.lib()
buf = torch.zeros(148 * 8 + 2 * 18 * 16, dtype=torch.int64, device='cuda')
ops.mlp_branch(t, w1, b1, w2, b2, cab=cab, cs=cs, out=out, tok_per_img=ntok // B)
torch.cuda.synchronize()
lib.srb_debug_set_phase_buffer.argtypes = [C.c_void_p]; lib.srb_debug_set_phase_buffer.restype = None
lib.srb_debug_set_phase_buffer(C.c_void_p(buf.data_ptr()))
ops.mlp_branch(t, w1, b1, w2, b2, cab=cab, cs=cs, out=out, tok_per_img=ntok // B)
torch.cuda.synchronize()
lib.srb_debug_set_phase_buffer(None)
tr = buf[148 * 8:].view(2, 18, 16).cpu()
t0 = int(tr[0, 0, 0])
AN = ['top', 'D1 ready', 'GELU done (a2)', 'after bar1', 'LN(i+1) done', 'D2 ready', 'EPI done', 'after bar2', 'copy-out done', 'loads issued']
MN = ['a1 ready', 'fc1 issued', 'a2 ready', 'fc2 issued']
ev = []
for ti in range(2):
    for w in (0, 5, 10, 15, 16):
        names = MN if w == 16 else AN
        for k, n in enumerate(names):
            v = int(tr[ti, w, k])
            if v:
                ev.append((v - t0, f'tile {3 + ti} ' + ('mma' if w == 16 else 'warp %2d' % w), n))
ev.sort()
for c, who, n in ev:
    print(f'{c:8d}  {who:16s} {n}')
