"""Timeline of hab_tc_kernel (build with EXTRA=-DSRB_HAB_TRACE): clock64 stamps of CTA 0, its tiles 3 and 4, per warp."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel
from superresolution_b200 import _lib as L
torch.manual_seed(0)
m = HybridSuperResolutionModel(output_size=512).cuda().eval()
x = torch.rand(16, 1, 128, 128, device='cuda')
lib = L.lib()
buf = torch.zeros(2 * 148 * 8 + 2 * 16 * 32, dtype=torch.int64, device='cuda')
with torch.no_grad():
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer.argtypes = [C.c_void_p]; lib.srb_debug_set_phase_buffer.restype = None
    lib.srb_debug_set_phase_buffer(C.c_void_p(buf.data_ptr()))
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer(None)
t = buf[2 * 148 * 8:].view(2, 16, 32).cpu()
t0 = int(t[0, 0, 0])
CN = ['top', 'qkv ready', 'qkv epi done', 'x landed', 'LN done', 'S h0 ready', 'P h0 done', 'S h1 ready', 'P h1 done',
      'O h0 ready', 'O h1 ready', 'ao arrive', 'proj ready', 'staged', 'before S wait', 'proj epi done']
IN = ['qkvw ready', 'S issued', 'P h0 ready', 'PV h0 issued', 'P h1 ready', 'PV h1 issued', 'ao ready', 'proj issued',
      'a1 ready', 'qkv issued']
ION = ['a1 ready', 'xn copied out', 'qkv x3 ready', 'fetch issued', 'x landed']
ev = []
for ti in range(2):
    for w in range(16):
        names = CN if w < 12 else ION if w == 12 else IN
        if w < 12 and w % 4:
            continue
        for k, n in enumerate(names):
            v = int(t[ti, w, k])
            if v and n:
                ev.append((v - t0, f'tile {3 + ti} ' + ('group %d' % (w // 4) if w < 12 else 'io' if w == 12 else 'issuer %d' % (w - 13)), n))
ev.sort()
for c, who, n in ev:
    print(f'{c:8d}  {who:18s} {n}')
