"""One profiled forward at the bench geometry (B=16, 128x128, balanced) for `ncu --profile-from-start off`:
warm-up forwards run outside the cudaProfilerStart/Stop window, exactly one forward (184 launches) inside it."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel
torch.manual_seed(0)
m = HybridSuperResolutionModel(output_size=512).cuda().eval()
g = torch.Generator().manual_seed(1)
x = (torch.randint(0, 65536, (16, 1, 128, 128), generator=g).float() / 65535.0).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
with torch.no_grad():
    for _ in range(2):
        m(x)
    flush.zero_()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    y = m(x)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
print('ok', float(y.mean()))
