"""HAB attention branch alone at the bench geometry, for ncu (-k regex:hab_tc) and quick timing."""
import os, sys, torch, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel
torch.manual_seed(0)
m = HybridSuperResolutionModel(output_size=512).cuda().eval()
x2 = torch.rand(16, 1, 256, 256, device='cuda')
with torch.no_grad():
    for _ in range(2):
        m.run(x2, stage_mask=2)
torch.cuda.synchronize()
