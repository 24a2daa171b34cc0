import sys, ctypes as C, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel
from superresolution_b200 import _lib as L
torch.manual_seed(0)
m = HybridSuperResolutionModel(output_size=512).cuda().eval()
x = torch.rand(16,1,128,128,device='cuda')
lib = L.lib()
buf = torch.zeros(2*148*8, dtype=torch.int64, device='cuda')
with torch.no_grad():
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer.argtypes=[C.c_void_p]; lib.srb_debug_set_phase_buffer.restype=None
    lib.srb_debug_set_phase_buffer(C.c_void_p(buf.data_ptr()))
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer(None)
names=['wait_qkv','qkv_epi','S_issue','softmax+PV','LN(i+1)','O_norm','proj_wait+epi']
for w, what in ((0, 'warp 0 (group leader: issues S / PV)'), (1, 'warp 1')):
    p = buf.view(2,148,8)[w].double().mean(0)
    tot=p[:7].sum()
    print(what)
    for n,v in zip(names,p[:7]): print(f'  {n:14s} {v.item()/1e3:9.1f} kcyc {100*v.item()/tot.item():5.1f}%  {v.item()/(8192/148):8.0f} cycles per tile')
    print('  total kcycles per CTA', tot.item()/1e3, 'tiles per CTA ~', 8192/148, 'cycles per 128-token tile', tot.item()/(8192/148))
