"""Graph-replay timing of stage 1 (RRDBNet), stage 2 (HAT) and the full forward at the bench geometry (B=16, 128x128)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200.architecture import HybridSuperResolutionModel
torch.manual_seed(0)
m = HybridSuperResolutionModel(output_size=512).cuda().eval()
x1 = torch.rand(16, 1, 128, 128, device='cuda')
x2 = torch.rand(16, 1, 256, 256, device='cuda')
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for name, x, mask in (('stage1', x1, 1), ('stage2', x2, 2), ('full', x1, 3)):
    with torch.no_grad():
        m.run(x, stage_mask=mask); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            m.run(x, stage_mask=mask)
            with torch.cuda.graph(g, stream=s):
                m.run(x, stage_mask=mask)
        ts = []
        for _ in range(6):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print(f'{name}: min {min(ts):.3f} ms  median {sorted(ts)[len(ts)//2]:.3f} ms')
