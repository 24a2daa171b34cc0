"""Optimizer half of the reference training step (``scripts/Modello_supporto.py:103-107,133-141``):

    nn.DataParallel reduce-add  ->  ``allreduce_gradients`` (NCCL all-reduce of ONE flat fp32 gradient buffer, in buckets)
    scaler.unscale_ + clip_grad_norm_(1.0) + AdamW.step  ->  ``FusedAdamW.step(grad_scale=, max_norm=, world=)``
                                                             (two libsrb kernels, no host synchronisation)

``FusedAdamW`` is a ``torch.optim.Optimizer``: it can stand where ``optim.AdamW(model.parameters(), lr, weight_decay=0)``
stands, works with ``CosineAnnealingLR`` (reads ``param_groups[0]['lr']`` every step) and with the stock
``GradScaler.unscale_/step`` + ``clip_grad_norm_`` sequence (plain ``step()``), and additionally offers the fused
form.  Parameters, gradients and both moments are flat fp32 buffers; every ``p.data`` / ``p.grad`` is a view into
them, and ``zero_grad`` zeroes the flat gradient instead of dropping the views.

The gradients come from the libsrb backward of both stages (training.py / training_hat.py) or from any other autograd
graph over the parameters; ``state_dict()`` carries the step count and both moment buffers (flat), so a run resumes
where ``torch.optim.AdamW`` would.  One deviation from ``GradScaler``: a fused step that the kernel skips because of an
inf/nan gradient still advances the bias-correction count (the skip is decided on the device, without a host
synchronisation); ``settle_skipped_step()`` reads the flag (one sync) and takes the count back.
"""
import torch
import torch.distributed as dist

from . import _lib as L


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, bf16_shadow=False):
        params = [p for p in params]
        if not params:
            raise ValueError('FusedAdamW: empty parameter list')
        if any(isinstance(p, dict) for p in params):
            raise ValueError('FusedAdamW: a single parameter group (as the reference uses)')
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        ps = self.param_groups[0]['params']
        dev = ps[0].device
        if dev.type != 'cuda':
            raise RuntimeError('FusedAdamW (libsrb) runs on CUDA only; no CPU fallback')
        if any(p.device != dev or p.dtype != torch.float32 for p in ps):
            raise ValueError('FusedAdamW: all parameters must be fp32 on one CUDA device')
        # every tensor starts on a 16-byte boundary of the flat buffers (float4 kernels); pads stay zero
        self._offsets, n = [], 0
        for p in ps:
            self._offsets.append(n)
            n += (p.numel() + 3) // 4 * 4
        self.numel = n
        self.flat_param = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.shadow = torch.zeros(n, dtype=torch.bfloat16, device=dev) if bf16_shadow else None
        self._info = torch.zeros(2, dtype=torch.float32, device=dev)
        self._scratch = torch.empty(L.lib().srb_grad_norm_scratch_bytes(), dtype=torch.uint8, device=dev)
        self._step = 0
        with torch.no_grad():
            for p, o in zip(ps, self._offsets):
                view = self.flat_param[o:o + p.numel()].view_as(p)
                view.copy_(p)
                p.data = view
                g = self.flat_grad[o:o + p.numel()].view_as(p)
                if p.grad is not None:
                    g.copy_(p.grad)
                p.grad = g

    def zero_grad(self, set_to_none=True):  # noqa: ARG002 - the views must survive: zero instead of dropping
        self.flat_grad.zero_()
        for p, o in zip(self.param_groups[0]['params'], self._offsets):
            if p.grad is None or p.grad.data_ptr() != self.flat_grad.data_ptr() + 4 * o:
                p.grad = self.flat_grad[o:o + p.numel()].view_as(p)

    def offset_of(self, param):
        """first element of ``param`` in the flat buffers (the all-reduce of a parameter range can start as soon as the
        last gradient of that range is final)"""
        for p, o in zip(self.param_groups[0]['params'], self._offsets):
            if p is param:
                return o
        raise KeyError('FusedAdamW.offset_of: not a parameter of this optimizer')

    def _gather_stray_grads(self):
        """``model.zero_grad()`` (set_to_none is torch's default) or ``p.grad = None`` drops the views; autograd then
        allocates fresh ``.grad`` tensors and the flat buffer would go stale.  Before the flat buffer is read, every
        gradient that no longer aliases it is copied in and re-pointed (a missing gradient counts as zero, like
        torch.optim.AdamW skipping a parameter without ``.grad`` does for that step's gradient)."""
        base = self.flat_grad.data_ptr()
        for p, o in zip(self.param_groups[0]['params'], self._offsets):
            gr = p.grad
            if gr is not None and gr.data_ptr() == base + 4 * o:
                continue
            view = self.flat_grad[o:o + p.numel()].view_as(p)
            if gr is None:
                view.zero_()
            else:
                view.copy_(gr)
            p.grad = view

    def grad_norm(self, grad_scale=1.0, world=1):
        """L2 norm of the (unscaled, rank-averaged) gradient as a 0-dim CUDA tensor (no sync)."""
        self._sumsq()
        return self._info[0].sqrt() / (float(grad_scale) * world)

    def _sumsq(self):
        self._gather_stray_grads()
        with torch.cuda.device(self.flat_grad.device):
            L.check(L.lib().srb_grad_sumsq(L.ptr(self.flat_grad), self.numel, L.ptr(self._info), L.ptr(self._scratch),
                                           L.current_stream()), 'grad_sumsq')

    @torch.no_grad()
    def step(self, closure=None, grad_scale=None, max_norm=None, world=1):
        """Plain ``step()`` = torch.optim.AdamW.step on gradients that are already unscaled / clipped.
        ``step(grad_scale=s, max_norm=c, world=w)`` folds GradScaler.unscale_ (1/s), the rank mean (1/w, after a
        SUM all-reduce), clip_grad_norm_(c) and the inf/nan skip into the same two kernels."""
        loss = closure() if closure is not None else None
        g = self.param_groups[0]
        fused = grad_scale is not None or max_norm is not None or world != 1
        inv_scale = 1.0 / (float(grad_scale if grad_scale is not None else 1.0) * world)
        if fused:
            self._sumsq()
        else:
            self._gather_stray_grads()
        self._step += 1
        with torch.cuda.device(self.flat_param.device):
            L.check(L.lib().srb_adamw_step(
                L.ptr(self.flat_param), L.ptr(self.flat_grad), L.ptr(self.exp_avg), L.ptr(self.exp_avg_sq), self.numel,
                float(g['lr']), float(g['betas'][0]), float(g['betas'][1]), float(g['eps']), float(g['weight_decay']),
                self._step, inv_scale, float(max_norm) if max_norm is not None else 0.0,
                L.ptr(self._info) if fused else None, L.ptr(self.shadow), L.current_stream()), 'adamw_step')
        # the kernel wrote through the views: bump the version counters so that consumers keyed on them (the model's
        # packed-weight cache, autograd's saved-tensor checks) see the update
        torch.autograd.graph.increment_version(g['params'])
        return loss

    def settle_skipped_step(self):
        """GradScaler parity for the bias correction: if the last fused step was skipped (inf/nan gradient), take the
        step count back.  Reads the device flag (one host synchronisation); returns True when a step was skipped."""
        skipped = bool(self._info[1].item() != 0.0)
        if skipped:
            self._step = max(0, self._step - 1)
            self._info[1].zero_()  # settled: a second call does not take another step back
        return skipped

    def state_dict(self):
        """torch layout ('state' / 'param_groups') with ONE state entry: the flat moment buffers and the step count."""
        sd = super().state_dict()
        sd['state'] = {0: {'step': self._step, 'exp_avg': self.exp_avg.clone(), 'exp_avg_sq': self.exp_avg_sq.clone()}}
        return sd

    def load_state_dict(self, state_dict):
        st = state_dict['state'][0] if 0 in state_dict['state'] else state_dict['state']['0']
        if st['exp_avg'].numel() != self.numel or st['exp_avg_sq'].numel() != self.numel:
            raise ValueError('FusedAdamW.load_state_dict: moment buffers of %d elements expected' % self.numel)
        for k, v in state_dict['param_groups'][0].items():
            if k != 'params':
                self.param_groups[0][k] = v
        self._step = int(st['step'])
        self.exp_avg.copy_(st['exp_avg'])
        self.exp_avg_sq.copy_(st['exp_avg_sq'])

    def found_inf(self):
        """1.0 if the last fused step saw an inf/nan gradient and was skipped (0-dim CUDA tensor; feed to the scaler)."""
        return self._info[1]


def bucket_ranges(numel, bucket_elems):
    """[start, end) element ranges covering a flat buffer in buckets of at most ``bucket_elems``."""
    if bucket_elems <= 0:
        raise ValueError('bucket_elems must be positive')
    return [(s, min(s + bucket_elems, numel)) for s in range(0, numel, bucket_elems)]


def allreduce_gradients_start(flat_grad, group=None, bucket_bytes=25 << 20, lo=0, hi=None):
    """Launch the SUM all-reduce of flat_grad[lo:hi] in ~25 MB buckets and return the pending work handles: the
    collectives run on the backend's own stream, so whatever the caller enqueues next (the stage-1 backward, which does
    not touch this range) overlaps them.  Finish with ``allreduce_gradients_wait``."""
    if not dist.is_available() or not dist.is_initialized():
        raise RuntimeError('allreduce_gradients: torch.distributed is not initialised')
    if dist.get_world_size(group) == 1:
        return []
    hi = flat_grad.numel() if hi is None else hi
    per = max(1, bucket_bytes // flat_grad.element_size())
    return [dist.all_reduce(flat_grad[lo + s:lo + e], op=dist.ReduceOp.SUM, group=group, async_op=True)
            for s, e in bucket_ranges(hi - lo, per)]


def allreduce_gradients_wait(works):
    for w in works:
        w.wait()


def allreduce_gradients(flat_grad, group=None, bucket_bytes=25 << 20):
    """SUM all-reduce of the flat gradient in ~25 MB buckets (all launched asynchronously, then waited), the
    replacement for nn.DataParallel's reduce-add (Modello_supporto.py:103-105).  The 1/world mean is NOT applied
    here: pass ``world=`` to FusedAdamW.step, which folds it into the update.  Backend: NCCL on CUDA tensors (NVLink /
    NVSwitch), gloo on CPU tensors (tests)."""
    allreduce_gradients_wait(allreduce_gradients_start(flat_grad, group, bucket_bytes))
    return flat_grad
