"""Drop-in for reference ``src/metrics.py``: ``Metrics`` / ``ssim_torch`` on the fused libsrb PSNR+SSIM kernel
(one pass over the pair instead of five 11x11 depthwise convs; src/metrics.py:5-35)."""
import torch

from . import _lib as L


def _psnr_ssim(p, t):
    """returns a 2-element CUDA tensor: [sum_b psnr_b, ssim_batch_mean * B]"""
    if p.dim() != 4 or p.shape[1] != 1:
        raise ValueError('expected [B,1,H,W] tensors')
    if p.device.type != 'cuda':
        raise RuntimeError('Metrics (libsrb) runs on CUDA only; no CPU fallback')
    p = p.detach().float().contiguous()
    t = t.detach().float().contiguous()
    B, _, H, W = p.shape
    lib = L.lib()
    out = torch.empty(2, dtype=torch.float32, device=p.device)
    scratch = torch.empty(lib.srb_psnr_ssim_scratch_bytes(B, H, W), dtype=torch.uint8, device=p.device)
    with torch.cuda.device(p.device):
        L.check(lib.srb_psnr_ssim(L.ptr(p), L.ptr(t), B, H, W, L.ptr(out), L.ptr(scratch), L.current_stream()),
                'psnr_ssim')
    return out


def ssim_torch(img1, img2, window_size=11):
    """batch-global mean SSIM of clamped inputs is what Metrics uses; this mirrors the reference function for
    already-clamped single-channel inputs (window_size fixed at 11)."""
    if window_size != 11:
        raise NotImplementedError('libsrb implements the 11x11 window of src/metrics.py')
    # the kernel saturates its inputs to [0,1] (what Metrics.update feeds it); the reference function itself does not
    # clamp, so anything else would silently differ: refuse it (one device->host sync; not on the Metrics path)
    for name, img in (('img1', img1), ('img2', img2)):
        if not bool(((img >= 0) & (img <= 1)).all()):
            raise ValueError(f'ssim_torch (libsrb): {name} has values outside [0,1] or NaN; clamp first '
                             f'(Metrics.update does) - the fused kernel saturates its inputs')
    return _psnr_ssim(img1, img2)[1] / img1.shape[0]


class Metrics:
    def __init__(self):
        self.reset()

    def reset(self):
        self.psnr = 0.0
        self.ssim = 0.0
        self.count = 0

    def update(self, p, t):
        r = _psnr_ssim(p, t).tolist()  # one device->host sync per call (the reference does two .item())
        self.psnr += r[0]
        self.ssim += r[1]
        self.count += p.size(0)

    def compute(self):
        if not self.count:
            return {'psnr': 0, 'ssim': 0}
        return {'psnr': self.psnr / self.count, 'ssim': self.ssim / self.count}
