"""Drop-in for reference ``src/losses.py``: ``CombinedLoss`` (star-weighted L1) on the libsrb kernel.

forward(pred, target) -> (loss, {'total': loss, 'weighted_l1': loss}) exactly like src/losses.py:9-21.  The loss is a
0-dim CUDA tensor; when ``pred`` requires grad the analytic gradient sign(p-t)*w/N produced by the same kernel pass
is attached through a tiny autograd.Function (one fused pass instead of ~6 elementwise kernels + a masked scatter).
"""
import torch
import torch.nn as nn

from . import _lib as L


def _scratch(dev):
    return torch.empty(L.lib().srb_loss_scratch_bytes(), dtype=torch.uint8, device=dev)


class _WeightedL1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        p = pred.detach().float().contiguous()
        t = target.detach().float().contiguous()
        if p.device.type != 'cuda':
            raise RuntimeError('CombinedLoss (libsrb) runs on CUDA only; no CPU fallback')
        loss = torch.empty((), dtype=torch.float32, device=p.device)
        grad = torch.empty_like(p) if pred.requires_grad else None
        with torch.cuda.device(p.device):
            L.check(L.lib().srb_weighted_l1(L.ptr(p), L.ptr(t), p.numel(), L.ptr(loss), L.ptr(grad),
                                            L.ptr(_scratch(p.device)), L.current_stream()), 'weighted_l1')
        ctx.save_for_backward(grad)
        ctx.in_dtype = pred.dtype
        return loss

    @staticmethod
    def backward(ctx, gout):
        (grad,) = ctx.saved_tensors
        return (grad * gout).to(ctx.in_dtype), None


class CombinedLoss(nn.Module):
    def __init__(self, l1_w=1.0, perceptual_w=0.0, astro_w=1.0):  # ctor args ignored, like the reference (:6-7)
        super().__init__()

    def forward(self, pred, target):
        loss = _WeightedL1.apply(pred, target)
        return loss, {'total': loss, 'weighted_l1': loss}


# ---------------------------------------------------------------------------------------------------------------------
# Stale variant: the Charbonnier and "astro" terms of reference Backup/src/losses.py:6-15,37-48 on one
# libsrb pass.  That file's third term is a VGG19 perceptual loss on downloaded ImageNet weights: not built (no network,
# and not on the live path) - asking for it raises.
# ---------------------------------------------------------------------------------------------------------------------
class _CharbonnierAstro(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target, l1_w, astro_w):
        p = pred.detach().float().contiguous()
        t = target.detach().float().contiguous()
        if p.device.type != 'cuda':
            raise RuntimeError('CharbonnierLoss / AstroCombinedLoss (libsrb) run on CUDA only; no CPU fallback')
        out3 = torch.empty(3, dtype=torch.float32, device=p.device)
        grad = torch.empty_like(p) if pred.requires_grad else None
        with torch.cuda.device(p.device):
            L.check(L.lib().srb_charbonnier_astro(L.ptr(p), L.ptr(t), p.numel(), float(l1_w), float(astro_w), L.ptr(out3),
                                                  L.ptr(grad), L.ptr(_scratch(p.device)), L.current_stream()),
                    'charbonnier_astro')
        ctx.save_for_backward(grad)
        ctx.in_dtype = pred.dtype
        ctx.mark_non_differentiable(out3)
        total = l1_w * out3[0] + astro_w * out3[2]
        return total, out3

    @staticmethod
    def backward(ctx, gout, _g3):
        (grad,) = ctx.saved_tensors
        return (grad * gout).to(ctx.in_dtype), None, None, None


class CharbonnierLoss(nn.Module):
    """mean(sqrt((x - y)^2 + eps)) with the reference's eps = 1e-6 (Backup/src/losses.py:6-15)."""

    def __init__(self, eps=1e-6):
        super().__init__()
        if eps != 1e-6:
            raise NotImplementedError('libsrb implements the eps = 1e-6 of Backup/src/losses.py')

    def forward(self, x, y):
        return _CharbonnierAstro.apply(x, y, 1.0, 0.0)[0]


class AstroCombinedLoss(nn.Module):
    """``CombinedLoss`` of Backup/src/losses.py:17-70 without its VGG term: total = l1_w * char + astro_w * astro and the
    same log dict ('char', 'l1_raw', 'astro'; 'perceptual' = 0)."""

    def __init__(self, l1_w=1.0, perceptual_w=0.0, astro_w=0.05):
        super().__init__()
        if perceptual_w != 0.0:
            raise NotImplementedError('the VGG19 perceptual term (downloaded ImageNet weights) is not built: pass '
                                      'perceptual_w=0.0')
        self.weights = (l1_w, perceptual_w, astro_w)

    def forward(self, pred, target):
        total, out3 = _CharbonnierAstro.apply(pred, target, self.weights[0], self.weights[2])
        return total, {'char': out3[0], 'l1_raw': out3[1], 'astro': out3[2], 'perceptual': out3.new_zeros(())}
