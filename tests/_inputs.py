"""Seeded synthetic inputs shared by the golden generator and the tests (same code as tests/golden/make_golden.py)."""
import numpy as np
import torch


def model_inputs():
    g = torch.Generator().manual_seed(1)
    x32 = torch.randint(0, 65536, (1, 1, 32, 32), generator=g).float() / 65535.0
    g = torch.Generator().manual_seed(2)
    x64 = torch.randint(0, 65536, (1, 1, 64, 64), generator=g).float() / 65535.0
    return x32, x64


def aux_inputs():
    g = torch.Generator().manual_seed(11)
    t = torch.rand(2, 1, 96, 80, generator=g) * 0.03
    t[0, 0, 10:14, 20:24] = 0.9
    p = (t + 0.05 * torch.randn(2, 1, 96, 80, generator=g))
    rng = np.random.default_rng(3)
    raw = np.expm1(rng.normal(1.0, 1.2, size=(64, 48))).astype(np.float32)
    raw[rng.random(raw.shape) < 0.01] *= -1.0
    raw[3, 5] = np.nan
    raw[7, 9] = np.inf
    raw[8, 9] = -np.inf
    raw[0, 0] = 0.0
    return p, t, raw
