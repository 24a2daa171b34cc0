import sys, ctypes as C, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel
from superresolution_b200 import _lib as L
torch.manual_seed(0)
m = HybridSuperResolutionModel(output_size=512).cuda().eval()
x = torch.rand(16,1,128,128,device='cuda')
lib = L.lib()
buf = torch.zeros(148*8, dtype=torch.int64, device='cuda')
with torch.no_grad():
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer.argtypes=[C.c_void_p]; lib.srb_debug_set_phase_buffer.restype=None
    lib.srb_debug_set_phase_buffer(C.c_void_p(buf.data_ptr()))
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer(None)
p = buf.view(148,8).double().mean(0)
names=['layernorm','wait_qkv','qkv_epi','qkv_barrier','attention','wait_proj','proj_epi']
tot=p[:7].sum()
for n,v in zip(names,p[:7]): print(f'{n:14s} {v.item()/1e3:9.1f} kcyc {100*v.item()/tot.item():5.1f}%')
print('total kcycles per CTA', tot.item()/1e3, 'tiles per CTA ~', 8192/148, 'cycles per 128-token tile', tot.item()/(8192/148))
