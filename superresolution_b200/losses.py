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
