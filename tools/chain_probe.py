"""Per-role cycle breakdown of the RRDB chain kernel (last chain launch of a stage-1-only forward, B=16, 128x128)."""
import sys, ctypes as C, torch, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel
from superresolution_b200 import _lib as L
torch.manual_seed(0)
m = HybridSuperResolutionModel(output_size=512).cuda().eval()
x = torch.rand(16, 1, 128, 128, device='cuda')
lib = L.lib()
buf = torch.zeros(148 * 16, dtype=torch.int64, device='cuda')
with torch.no_grad():
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer.argtypes = [C.c_void_p]; lib.srb_debug_set_phase_buffer.restype = None
    lib.srb_debug_set_phase_buffer(C.c_void_p(buf.data_ptr()))
    m(x); torch.cuda.synchronize()
    lib.srb_debug_set_phase_buffer(None)
p = buf.view(148, 16)[:128].double().mean(0) / 15.0  # per layer
names = ['producer: grid-barrier wait', 'producer: total', 'mma: wait patch', 'mma: wait weights', 'mma: wait acc_empty',
         'mma: total', 'epi: wait acc_full', 'epi: epilogue', 'epi: store-complete + arrive']
for n, v in zip(names, p):
    print(f'{n:32s} {v.item():9.0f} cyc/layer  {v.item() / 1.9e3:6.2f} us')
