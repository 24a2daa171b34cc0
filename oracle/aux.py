"""ORACLE (test infrastructure): the HBM-bound neighbours of the model forward, restated from the reference's own
code (these files ARE in /root/reference, so they are behavioural 1:1 restatements):

* ``combined_loss``      — reference src/losses.py:9-21
* ``ssim_torch``/Metrics — reference src/metrics.py:5-35
* ``u16_to_unit``        — reference src/dataset.py:31-34
* ``log_normalize_u16``  — reference scripts/Dataset_step4_normalization.py:161-179
* ``to_u16``             — reference scripts/Modello_5.py:29-33 (clip, *65535, truncating cast)

``tests/golden/make_golden.py`` checks these against the reference functions imported from /root/reference.
"""
from math import exp

import numpy as np
import torch
import torch.nn.functional as F


def combined_loss(pred, target):
    diff = torch.abs(pred - target)
    weight = torch.ones_like(diff)
    weight[target > 0.02] = 500.0
    return torch.mean(diff * weight)


def charbonnier_loss(x, y, eps=1e-6):
    """CharbonnierLoss.forward of the stale variant, reference Backup/src/losses.py:6-15."""
    diff = x - y
    return torch.mean(torch.sqrt(diff * diff + eps))


def astro_loss_terms(pred, target):
    """The non-VGG terms of the stale CombinedLoss, reference Backup/src/losses.py:37-48: (char, l1_raw, astro).
    The perceptual (VGG19) term needs downloaded ImageNet weights and stays out."""
    char = charbonnier_loss(pred, target)
    l1_raw = F.l1_loss(pred, target)
    diff = torch.abs(pred - target)
    weight_map = 1.0 + 5.0 * target
    astro = torch.mean(torch.sqrt(diff * diff + 1e-6) * weight_map)
    return char, l1_raw, astro


def ssim_torch(img1, img2, window_size=11):
    c = img1.size(1)
    g = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(window_size)])
    g = g / g.sum()
    win = g.unsqueeze(1).mm(g.unsqueeze(0)).unsqueeze(0).unsqueeze(0).expand(c, 1, window_size, window_size).type_as(img1)
    mu1, mu2 = F.conv2d(img1, win, groups=c), F.conv2d(img2, win, groups=c)
    s1 = F.conv2d(img1 * img1, win, groups=c) - mu1.pow(2)
    s2 = F.conv2d(img2 * img2, win, groups=c) - mu2.pow(2)
    s12 = F.conv2d(img1 * img2, win, groups=c) - mu1 * mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1 * mu2 + c1) * (2 * s12 + c2)) / ((mu1.pow(2) + mu2.pow(2) + c1) * (s1 + s2 + c2))).mean()


class Metrics:
    def __init__(self):
        self.reset()

    def reset(self):
        self.psnr = 0.0
        self.ssim = 0.0
        self.count = 0

    def update(self, p, t):
        p = p.clamp(0, 1)
        t = t.clamp(0, 1)
        mse = F.mse_loss(p, t, reduction='none').mean(dim=[1, 2, 3])
        self.psnr += (10 * torch.log10(1.0 / (mse + 1e-8))).sum().item()
        self.ssim += ssim_torch(p, t).item() * p.size(0)
        self.count += p.size(0)

    def compute(self):
        if not self.count:
            return {'psnr': 0, 'ssim': 0}
        return {'psnr': self.psnr / self.count, 'ssim': self.ssim / self.count}


def u16_to_unit(arr_u16):
    return np.asarray(arr_u16, dtype=np.float32) / 65535.0


def log_normalize_u16(data, gmin, gmax):
    """float32 image -> uint16 (truncating cast), float32 arithmetic throughout like the reference."""
    d = np.nan_to_num(np.asarray(data, dtype=np.float32))
    d = np.log1p(np.maximum(d, 0))
    gmin = np.float32(gmin)
    gmax = np.float32(gmax)
    d = (d - gmin) / (gmax - gmin + 1e-8)
    d = np.clip(d, 0, 1)
    return (d * 65535).astype(np.uint16)


def to_u16(t):
    return (np.clip(np.asarray(t, dtype=np.float32), 0, 1) * 65535).astype(np.uint16)


def robust_stats(images, black_clip_percentile=3.0, num_samples_per_img=None, batch_size=32):
    """calculate_robust_stats of scripts/Dataset_step4_normalization.py:55-91 on in-memory float32 images (the FITS read
    of :42-51 — nan_to_num(nan, +-inf -> 0), cast to float32 — is applied here).  Batches of ``batch_size`` images are
    stacked like the DataLoader does (images of one call must share a shape).  ``num_samples_per_img=None`` keeps EVERY
    valid pixel instead of the reference's unseeded ``torch.randperm`` subsample of 4000 per image (:73-75): that is
    the population value the reference's estimate scatters around, and the one the CUDA path computes exactly."""
    import torch
    sampled = []
    for s in range(0, len(images), batch_size):
        batch = torch.stack([torch.from_numpy(np.nan_to_num(np.asarray(im), nan=0.0, posinf=0.0, neginf=0.0)
                                              .astype(np.float32)) for im in images[s:s + batch_size]])
        batch = torch.log1p(torch.maximum(batch, torch.tensor(0.0)))
        flat = batch.view(-1)
        valid = flat[flat > 1e-5]
        if valid.numel() > 0:
            if num_samples_per_img is not None:
                take = min(valid.numel(), num_samples_per_img * batch.shape[0])
                valid = valid[torch.randperm(valid.numel())[:take]]
            sampled.append(valid.numpy())
    if not sampled:
        return 0.0, 1.0
    full = np.concatenate(sampled)
    return np.percentile(full, black_clip_percentile), np.percentile(full, 99.99)
