"""Drop-in for the reference's ``src/architecture.py``: same class names, constructor, ``forward(x)`` and state-dict
keys (reference src/architecture.py:9-82; RRDBNet / HAT layouts per SURVEY.md App. A), but the forward pass runs in
libsrb.so (hand-written sm_100a kernels behind the C ABI of include/srb.h).

The nn.Module tree below only *holds parameters* under the reference's names so that ``state_dict()``,
``load_state_dict()``, ``.parameters()``, ``.to()`` and ``nn.DataParallel(...).module`` behave as in the reference.
There is no PyTorch / CPU fallback: without a CUDA device and libsrb.so the forward raises.
"""
import ctypes as C

import torch
import torch.nn as nn

from . import _lib as L

_BINOMIAL = {3: [1, 2, 1], 5: [1, 4, 6, 4, 1], 7: [1, 6, 15, 20, 15, 6, 1]}
_DIVISOR = {3: 16.0, 5: 256.0, 7: 1600.0}  # the reference's literal divisors (7x7 sums to 2.56, :13)


def named_params(module):
    """``module.named_parameters()`` that also works on ``nn.DataParallel`` replicas: ``replicate()`` empties
    ``_parameters`` and keeps the broadcast copies (non-leaf, differentiable back to the originals) as plain attributes
    listed in ``_former_parameters`` (reference scripts/Modello_supporto.py:103-105 wraps the model in DataParallel)."""
    out = []
    for mname, mod in module.named_modules():
        replica = getattr(mod, '_is_replica', False) and hasattr(mod, '_former_parameters')
        src = mod._former_parameters if replica else mod._parameters
        for k, p in src.items():
            if p is not None:
                out.append(((mname + '.' if mname else '') + k, p))
    return out


def _conv(cin, cout, k=3):
    return nn.Conv2d(cin, cout, k, 1, k // 2)


def _no_direct_forward(self, *a, **k):
    raise RuntimeError('this sub-module only holds parameters; call HybridSuperResolutionModel.forward '
                       '(the forward pass runs in libsrb.so)')


class _ParamHolder(nn.Module):
    forward = _no_direct_forward


# ----------------------------------------------------------------------------------------------------------------
# stage 1: RRDBNet parameter tree (keys: conv_first, body.{i}.rdb{1,2,3}.conv{1..5}, conv_body, conv_up1, conv_up2,
# conv_hr, conv_last) — SURVEY.md App. A.1
# ----------------------------------------------------------------------------------------------------------------
class ResidualDenseBlock(_ParamHolder):
    def __init__(self, num_feat=64, num_grow_ch=32):
        super().__init__()
        for k in range(1, 6):
            conv = _conv(num_feat + (k - 1) * num_grow_ch, num_feat if k == 5 else num_grow_ch)
            nn.init.kaiming_normal_(conv.weight)  # default_init_weights(..., scale=0.1)
            conv.weight.data *= 0.1
            conv.bias.data.zero_()
            setattr(self, f'conv{k}', conv)


class RRDB(_ParamHolder):
    def __init__(self, num_feat, num_grow_ch=32):
        super().__init__()
        self.rdb1 = ResidualDenseBlock(num_feat, num_grow_ch)
        self.rdb2 = ResidualDenseBlock(num_feat, num_grow_ch)
        self.rdb3 = ResidualDenseBlock(num_feat, num_grow_ch)


class RRDBNet(_ParamHolder):
    def __init__(self, num_in_ch=1, num_out_ch=1, scale=2, num_feat=64, num_block=23, num_grow_ch=32):
        super().__init__()
        if (num_in_ch, num_out_ch, scale, num_feat, num_block, num_grow_ch) != (1, 1, 2, 64, 23, 32):
            raise NotImplementedError('libsrb implements the configuration of reference src/architecture.py:38-43')
        self.scale = scale
        self.conv_first = _conv(num_in_ch * 4, num_feat)
        self.body = nn.Sequential(*[RRDB(num_feat, num_grow_ch) for _ in range(num_block)])
        self.conv_body = _conv(num_feat, num_feat)
        self.conv_up1 = _conv(num_feat, num_feat)
        self.conv_up2 = _conv(num_feat, num_feat)
        self.conv_hr = _conv(num_feat, num_feat)
        self.conv_last = _conv(num_feat, num_out_ch)


# ----------------------------------------------------------------------------------------------------------------
# stage 2: HAT parameter tree — SURVEY.md App. A.2 / A.3
# ----------------------------------------------------------------------------------------------------------------
def _trunc_normal_(t, std=.02):
    return nn.init.trunc_normal_(t, mean=0., std=std, a=-2., b=2.)


class Mlp(_ParamHolder):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)


class WindowAttention(_ParamHolder):
    def __init__(self, dim, window_size, num_heads):
        super().__init__()
        self.relative_position_bias_table = nn.Parameter(torch.zeros((2 * window_size - 1) ** 2, num_heads))
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.proj = nn.Linear(dim, dim)
        _trunc_normal_(self.relative_position_bias_table)


class ChannelAttention(_ParamHolder):
    def __init__(self, num_feat, squeeze_factor):
        super().__init__()
        # indices 1 and 3 carry parameters (0 = AdaptiveAvgPool2d, 2 = ReLU, 4 = Sigmoid)
        self.attention = nn.Sequential(nn.Identity(), nn.Conv2d(num_feat, num_feat // squeeze_factor, 1),
                                       nn.Identity(), nn.Conv2d(num_feat // squeeze_factor, num_feat, 1),
                                       nn.Identity())


class CAB(_ParamHolder):
    def __init__(self, num_feat, compress_ratio, squeeze_factor):
        super().__init__()
        self.cab = nn.Sequential(_conv(num_feat, num_feat // compress_ratio), nn.Identity(),
                                 _conv(num_feat // compress_ratio, num_feat),
                                 ChannelAttention(num_feat, squeeze_factor))


class HAB(_ParamHolder):
    def __init__(self, dim, num_heads, window_size, compress_ratio, squeeze_factor, mlp_ratio):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn = WindowAttention(dim, window_size, num_heads)
        self.conv_block = CAB(dim, compress_ratio, squeeze_factor)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = Mlp(dim, int(dim * mlp_ratio))


class OCAB(_ParamHolder):
    def __init__(self, dim, window_size, overlap_ratio, num_heads, mlp_ratio):
        super().__init__()
        ows = int(window_size * overlap_ratio) + window_size
        self.norm1 = nn.LayerNorm(dim)
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.relative_position_bias_table = nn.Parameter(torch.zeros((window_size + ows - 1) ** 2, num_heads))
        _trunc_normal_(self.relative_position_bias_table)
        self.proj = nn.Linear(dim, dim)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = Mlp(dim, int(dim * mlp_ratio))


class AttenBlocks(_ParamHolder):
    def __init__(self, dim, depth, num_heads, window_size, compress_ratio, squeeze_factor, overlap_ratio, mlp_ratio):
        super().__init__()
        self.blocks = nn.ModuleList([HAB(dim, num_heads, window_size, compress_ratio, squeeze_factor, mlp_ratio)
                                     for _ in range(depth)])
        self.overlap_attn = OCAB(dim, window_size, overlap_ratio, num_heads, mlp_ratio)


class RHAG(_ParamHolder):
    def __init__(self, dim, depth, num_heads, window_size, compress_ratio, squeeze_factor, overlap_ratio, mlp_ratio):
        super().__init__()
        self.residual_group = AttenBlocks(dim, depth, num_heads, window_size, compress_ratio, squeeze_factor,
                                          overlap_ratio, mlp_ratio)
        self.conv = _conv(dim, dim)


class PatchEmbed(_ParamHolder):
    def __init__(self, dim):
        super().__init__()
        self.norm = nn.LayerNorm(dim)


def _rpi_sa(ws):
    c = torch.stack(torch.meshgrid([torch.arange(ws), torch.arange(ws)], indexing='ij')).flatten(1)
    rel = (c[:, :, None] - c[:, None, :]).permute(1, 2, 0).contiguous()
    rel[:, :, 0] += ws - 1
    rel[:, :, 1] += ws - 1
    rel[:, :, 0] *= 2 * ws - 1
    return rel.sum(-1)


def _rpi_oca(ws, overlap_ratio):
    wse = ws + int(overlap_ratio * ws)
    co = torch.stack(torch.meshgrid([torch.arange(ws), torch.arange(ws)], indexing='ij')).flatten(1)
    ce = torch.stack(torch.meshgrid([torch.arange(wse), torch.arange(wse)], indexing='ij')).flatten(1)
    rel = (ce[:, None, :] - co[:, :, None]).permute(1, 2, 0).contiguous()
    rel[:, :, 0] += ws - wse + 1
    rel[:, :, 1] += ws - wse + 1
    rel[:, :, 0] *= ws + wse - 1
    return rel.sum(-1)


class HAT(_ParamHolder):
    def __init__(self, img_size=64, patch_size=1, in_chans=1, embed_dim=96, depths=(6, 6, 6, 6),
                 num_heads=(6, 6, 6, 6), window_size=8, compress_ratio=3, squeeze_factor=30, conv_scale=0.01,
                 overlap_ratio=0.5, mlp_ratio=2., qkv_bias=True, upscale=2, img_range=1., upsampler='pixelshuffle',
                 resi_connection='1conv'):
        super().__init__()
        cfg = (patch_size, in_chans, embed_dim, tuple(depths), tuple(num_heads), window_size, compress_ratio,
               squeeze_factor, conv_scale, overlap_ratio, mlp_ratio, qkv_bias, upscale, img_range, upsampler,
               resi_connection)
        if cfg != (1, 1, 96, (6,) * 4, (6,) * 4, 8, 3, 30, 0.01, 0.5, 2., True, 2, 1., 'pixelshuffle', '1conv'):
            raise NotImplementedError('libsrb implements the configuration of reference src/architecture.py:51-61')
        self.register_buffer('relative_position_index_SA', _rpi_sa(window_size))
        self.register_buffer('relative_position_index_OCA', _rpi_oca(window_size, overlap_ratio))
        self.conv_first = _conv(in_chans, embed_dim)
        self.patch_embed = PatchEmbed(embed_dim)
        self.layers = nn.ModuleList([RHAG(embed_dim, d, nh, window_size, compress_ratio, squeeze_factor,
                                          overlap_ratio, mlp_ratio) for d, nh in zip(depths, num_heads)])
        self.norm = nn.LayerNorm(embed_dim)
        self.conv_after_body = _conv(embed_dim, embed_dim)
        self.conv_before_upsample = nn.Sequential(_conv(embed_dim, 64), nn.Identity())
        self.upsample = nn.Sequential(_conv(64, 4 * 64), nn.Identity())
        self.conv_last = _conv(64, in_chans)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                _trunc_normal_(m.weight)
                nn.init.constant_(m.bias, 0)


# ----------------------------------------------------------------------------------------------------------------
# the reference's own classes (src/architecture.py)
# ----------------------------------------------------------------------------------------------------------------
class AntiCheckerboardLayer(nn.Module):
    """reference src/architecture.py:9-27.  Persistent buffer ``weight`` [1,1,k,k]; inside the hybrid model the blur
    is fused into the libsrb forward (mode taken from the kernel size), so this module only carries the buffer."""

    def __init__(self, mode='balanced'):
        super().__init__()
        k = 7 if mode == 'strong' else 5 if mode == 'balanced' else 3
        self.mode = 'strong' if k == 7 else 'balanced' if k == 5 else 'light'
        row = torch.tensor(_BINOMIAL[k], dtype=torch.float32)
        self.register_buffer('weight', (torch.outer(row, row) / _DIVISOR[k]).unsqueeze(0).unsqueeze(0))
        self.padding = k // 2

    forward = _no_direct_forward


class HybridSuperResolutionModel(nn.Module):
    """Drop-in for reference src/architecture.py:30-82 (ctor signature, attributes, state-dict keys, forward(x))."""

    def __init__(self, output_size=512, smoothing='balanced', device='cpu'):
        super().__init__()
        self.output_size = output_size  # `device` is accepted and ignored, like the live reference (:31)
        self.stage1 = RRDBNet(num_in_ch=1, num_out_ch=1, num_feat=64, num_block=23, num_grow_ch=32, scale=2)
        self.stage2 = HAT(img_size=64, patch_size=1, in_chans=1, embed_dim=96, depths=[6, 6, 6, 6],
                          num_heads=[6, 6, 6, 6], window_size=8, compress_ratio=3, squeeze_factor=30,
                          conv_scale=0.01, overlap_ratio=0.5, mlp_ratio=2., qkv_bias=True, upscale=2, img_range=1.,
                          upsampler='pixelshuffle', resi_connection='1conv')
        self.has_stage2 = True  # never silently dropped (the reference swallows HAT ctor errors, :64-65)
        if smoothing != 'none':
            self.s1 = AntiCheckerboardLayer(smoothing)
            self.s2 = AntiCheckerboardLayer(smoothing)
            self.sf = AntiCheckerboardLayer('light')
        else:
            self.s1 = self.s2 = self.sf = nn.Identity()
        self._smoothing = L.SMOOTHING[self.s1.mode] if smoothing != 'none' else L.SMOOTHING['none']
        # engine state (never part of state_dict).  One entry per CUDA device: nn.DataParallel replicas share this dict
        # (replicate() copies __dict__ shallowly) and each replica thread only touches its own device's entry.
        self._names = None
        self._engine = {}

    # ---- engine --------------------------------------------------------------------------------------------
    def _param_list(self):
        lib = L.lib()
        if self._names is None:
            self._names = [lib.srb_param_name(i).decode() for i in range(lib.srb_param_count())]
        sd = dict(named_params(self))
        sd.update(dict(self.named_buffers()))
        return [sd[n] for n in self._names]

    def _dev_state(self, dev):
        st = self._engine.get(dev.index)
        if st is None:
            st = self._engine[dev.index] = {'packed': None, 'key': None, 'ws': {}}
        return st

    def pack_weights(self, force=False):
        """(Re)build the kernel-ready bf16 / fp32 weight image when parameters changed (tracked by tensor version
        counters and storage pointers).  Returns the packed device buffer."""
        lib = L.lib()
        tensors = self._param_list()
        key = tuple((t.data_ptr(), t._version) for t in tensors)
        dev = tensors[0].device
        if dev.type != 'cuda':
            raise RuntimeError('HybridSuperResolutionModel (libsrb) needs its parameters on a CUDA device; '
                               'there is no CPU path. Call .to("cuda").')
        st = self._dev_state(dev)
        if force or st['packed'] is None or key != st['key']:
            srcs = []
            for n, t in zip(self._names, tensors):
                want = torch.int64 if 'relative_position_index' in n else torch.float32
                srcs.append(t.detach().to(want).contiguous())
            ptrs = (C.c_void_p * len(srcs))(*[t.data_ptr() for t in srcs])
            if st['packed'] is None:
                st['packed'] = torch.empty(lib.srb_packed_bytes(), dtype=torch.uint8, device=dev)
            with torch.cuda.device(dev):
                L.check(lib.srb_pack_weights(ptrs, len(srcs), L.ptr(st['packed']), st['packed'].numel(),
                                             L.current_stream()), 'pack_weights')
            # srcs must outlive the enqueued copies: same stream ordering + caching allocator make this safe
            st['key'] = key
        return st['packed']

    def _workspace(self, B, H, W, dev):
        k = (B, H, W)
        cache = self._dev_state(dev)['ws']
        ws = cache.get(k)
        if ws is None:
            cache.clear()
            ws = torch.empty(L.lib().srb_workspace_bytes(B, H, W), dtype=torch.uint8, device=dev)
            cache[k] = ws
        return ws

    def run(self, x, stage_mask=3, out_dtype=torch.float32, out=None):
        """x: [B,1,H,W] float32 in [0,1] (or uint16 counts, divided by 65535 like src/dataset.py:34).
        stage_mask 1: RRDBNet only -> [B,1,2H,2W]; 2: HAT only -> [B,1,2H,2W]; 3: full model -> [B,1,4H,4W].
        out: optional preallocated result tensor (contiguous, right shape / dtype / device) instead of a fresh one."""
        lib = L.lib()
        if x.dim() != 4 or x.shape[1] != 1:
            raise ValueError(f'expected [B,1,H,W], got {tuple(x.shape)}')
        if x.device.type != 'cuda':
            raise RuntimeError('libsrb runs on CUDA only (no CPU fallback); move the input to the GPU')
        if x.dtype == torch.uint16:
            xd = L.SRB_U16
        elif x.dtype == torch.bfloat16:
            xd = L.SRB_BF16
        else:
            x = x.float()
            xd = L.SRB_F32
        x = x.contiguous()
        B, _, H, W = x.shape
        scale = 4 if stage_mask in (3, 7) else 2
        yd = L.SRB_U16 if out_dtype == torch.uint16 else L.SRB_F32
        with torch.cuda.device(x.device):
            packed = self.pack_weights()
            ws = self._workspace(B, H, W, x.device)
            ydt = torch.uint16 if yd == L.SRB_U16 else torch.float32
            if out is None:
                y = torch.empty(B, 1, scale * H, scale * W, dtype=ydt, device=x.device)
            else:
                if (tuple(out.shape) != (B, 1, scale * H, scale * W) or out.dtype != ydt or out.device != x.device
                        or not out.is_contiguous()):
                    raise ValueError('run(out=): expected a contiguous %s tensor of shape %s on %s'
                                     % (ydt, (B, 1, scale * H, scale * W), x.device))
                y = out
            L.check(lib.srb_forward(L.ptr(x), xd, L.ptr(y), yd, B, H, W, L.ptr(packed), L.ptr(ws), ws.numel(),
                                    self._smoothing, stage_mask, L.current_stream()), 'forward')
        return y

    PROFILE_CLASSES = ('conv3x3_tcgen05', 'linear_tcgen05', 'hab_attention_tcgen05', 'layernorm', 'other',
                       'mlp_tcgen05', 'ocab_attention', 'rrdb_chain_tcgen05')

    def profile(self, x, stage_mask=3):
        """bench helper: one forward with every launch bracketed by CUDA events on the launching stream.
        Returns {class: (ms, algorithmic_flops, launches)}; synchronises."""
        lib = L.lib()
        x = x.float().contiguous()
        B, _, H, W = x.shape
        scale = 4 if stage_mask == 3 else 2
        n = len(self.PROFILE_CLASSES)
        ms, fl, cnt = (C.c_double * n)(), (C.c_double * n)(), (C.c_int * n)()
        with torch.cuda.device(x.device):
            packed = self.pack_weights()
            ws = self._workspace(B, H, W, x.device)
            y = torch.empty(B, 1, scale * H, scale * W, dtype=torch.float32, device=x.device)
            L.check(lib.srb_forward_profile(L.ptr(x), L.SRB_F32, L.ptr(y), L.SRB_F32, B, H, W, L.ptr(packed),
                                            L.ptr(ws), ws.numel(), self._smoothing, stage_mask, L.current_stream(),
                                            ms, fl, cnt), 'forward_profile')
        return {k: (ms[i], fl[i], cnt[i]) for i, k in enumerate(self.PROFILE_CLASSES)}

    def launch_count(self, B, H, W, stage_mask=3):
        return L.lib().srb_forward_launch_count(B, H, W, self._smoothing, stage_mask)

    def forward(self, x):
        """src/architecture.py:74-82.  Inference (eval() or no_grad): one srb_forward call.  Training (train() mode with
        grad enabled): the differentiable libsrb path of training.hybrid_forward."""
        if torch.is_grad_enabled() and self.training and any(p.requires_grad for _, p in named_params(self)):
            from .training import hybrid_forward
            return hybrid_forward(self, x)
        if 4 * x.shape[-1] == self.output_size:
            return self.run(x, 3)
        # resize branch (:80-81): stage1 -> s1 -> stage2 -> s2, bicubic to output_size x output_size, then sf
        from . import ops
        y = ops.bicubic_resize(self.run(x, 7), self.output_size, self.output_size)
        return ops.blur(y, 'light') if self._smoothing else y
