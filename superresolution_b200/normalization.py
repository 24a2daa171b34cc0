"""LOG / 16-bit normalisation of reference ``scripts/Dataset_step4_normalization.py:161-179`` and the u16 <-> [0,1]
edges (``src/dataset.py:31-34``, ``scripts/Modello_5.py:29-33``) on the libsrb kernels."""
import numpy as np
import torch

from . import _lib as L


def log_normalize_u16(data, gmin, gmax):
    """data: float32 CUDA tensor (any shape) -> uint16 tensor of the same shape.
    u16(clip((log1p(max(nan_to_num(x), 0)) - gmin) / (gmax - gmin + 1e-8), 0, 1) * 65535), float32 arithmetic."""
    if data.device.type != 'cuda':
        raise RuntimeError('log_normalize_u16 (libsrb) runs on CUDA only; no CPU fallback')
    x = data.float().contiguous()
    out = torch.empty(x.shape, dtype=torch.uint16, device=x.device)
    gmin32 = np.float32(gmin)
    denom = np.float32(np.float32(gmax) - gmin32 + 1e-8)  # same float32 rounding as the numpy expression
    with torch.cuda.device(x.device):
        L.check(L.lib().srb_log_normalize_u16(L.ptr(x), L.ptr(out), x.numel(), float(gmin32), float(denom),
                                              L.current_stream()), 'log_normalize_u16')
    return out
