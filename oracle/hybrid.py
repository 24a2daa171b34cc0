"""ORACLE (test infrastructure): restatement of reference ``src/architecture.py:9-82`` on top of the restated
RRDBNet / HAT.  Same constructor, forward and state-dict keys as the reference class.  Parity unpinned (no
reference tests exist); RRDBNet/HAT come from ``oracle/rrdbnet.py`` / ``oracle/hat.py``.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .rrdbnet import RRDBNet
from .hat import HAT

_BINOMIAL = {3: [1, 2, 1], 5: [1, 4, 6, 4, 1], 7: [1, 6, 15, 20, 15, 6, 1]}
_DIVISOR = {3: 16.0, 5: 256.0, 7: 1600.0}  # the reference's literal divisors (7x7 sums to 2.56, :13)


class AntiCheckerboardLayer(nn.Module):
    """reference src/architecture.py:9-27 — fixed binomial depthwise blur, zero padding, buffer ``weight``."""

    def __init__(self, mode='balanced'):
        super().__init__()
        k = 7 if mode == 'strong' else 5 if mode == 'balanced' else 3
        row = torch.tensor(_BINOMIAL[k], dtype=torch.float32)
        # the reference divides the integer outer product by 1600 / 256 / 16
        kernel = torch.outer(row, row) / _DIVISOR[k]
        self.register_buffer('weight', kernel.unsqueeze(0).unsqueeze(0))
        self.padding = k // 2

    def forward(self, x):
        return F.conv2d(x, self.weight.expand(x.shape[1], -1, -1, -1), padding=self.padding, groups=x.shape[1])


class HybridSuperResolutionModel(nn.Module):
    """reference src/architecture.py:30-82."""

    def __init__(self, output_size=512, smoothing='balanced', device='cpu'):
        super().__init__()
        self.output_size = output_size
        self.stage1 = RRDBNet(num_in_ch=1, num_out_ch=1, num_feat=64, num_block=23, num_grow_ch=32, scale=2)
        self.stage2 = HAT(img_size=64, patch_size=1, in_chans=1, embed_dim=96, depths=[6, 6, 6, 6],
                          num_heads=[6, 6, 6, 6], window_size=8, compress_ratio=3, squeeze_factor=30,
                          conv_scale=0.01, overlap_ratio=0.5, mlp_ratio=2., qkv_bias=True, upscale=2,
                          img_range=1., upsampler='pixelshuffle', resi_connection='1conv')
        self.has_stage2 = True
        if smoothing != 'none':
            self.s1 = AntiCheckerboardLayer(smoothing)
            self.s2 = AntiCheckerboardLayer(smoothing)
            self.sf = AntiCheckerboardLayer('light')
        else:
            self.s1 = self.s2 = self.sf = nn.Identity()

    def forward(self, x):
        x = self.stage1(x)
        x = self.s1(x)
        if self.has_stage2:
            x = self.stage2(x)
            x = self.s2(x)
        if x.shape[-1] != self.output_size:
            x = F.interpolate(x, size=(self.output_size, self.output_size), mode='bicubic', align_corners=False)
        return self.sf(x)
