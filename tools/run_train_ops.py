"""The training-path operators of one HAB alone at the config-4 geometry (16 x 256 x 256 tokens): the TRAIN-epilogue
linears / convs and every weight-gradient shape, L2 flushed between launches, with the HBM bytes each must move."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from superresolution_b200 import ops, _lib as L

B, H, W = 16, 256, 256
dev = 'cuda'
torch.manual_seed(0)
mk = lambda c: torch.randn(B, H, W, c, device=dev).to(torch.bfloat16)
x96, y96, z96, x192, p192, x32, x288 = mk(96), mk(96), mk(96), mk(192), mk(192), mk(32), mk(288)
alpha = torch.ones(B, device=dev)
cs = torch.full((B, 96), 0.01, device=dev)
wl = lambda co, ci: ops.pack_conv_weight(torch.randn(co, ci, device=dev) / ci ** 0.5)
wc = lambda co, ci: ops.pack_conv_weight(torch.randn(co, ci, 3, 3, device=dev) / (9 * ci) ** 0.5)
wd = lambda co, ci: ops.pack_conv_weight_dgrad(torch.randn(co, ci, device=dev) / ci ** 0.5)
wdc = lambda co, ci: ops.pack_conv_weight_dgrad(torch.randn(co, ci, 3, 3, device=dev) / (9 * ci) ** 0.5)
bias = lambda c: torch.zeros(c, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
npx = B * H * W


def timed(name, fn, ch):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    tot, n = 0.0, 5
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    us = tot / n * 1e3
    nbytes = npx * ch * 2
    print(f'{name:44s} {us:8.1f} us   {nbytes / 1e6:7.0f} MB  {nbytes / us / 1e3:6.0f} GB/s')


w_fc1, w_fc2, w_proj, w_qkv = wl(192, 96), wl(96, 192), wl(96, 96), wl(288, 96)
o96, o192, o288, o32 = torch.empty_like(x96), torch.empty_like(x192), torch.empty_like(x288), torch.empty_like(x32)
pre192, pre32 = torch.empty_like(x192), torch.empty_like(x32)
timed('qkv fwd (plain)', lambda: ops.conv_nhwc(x96, 96, w_qkv, bias(288), out=o288), 96 + 288)
timed('fc1 fwd  (GELU + pre_out)', lambda: ops.conv_nhwc_train(x96, 96, w_fc1, bias(192), out=o192, act=L.ACT_GELU, pre_out=pre192), 96 + 384)
timed('fc2 fwd  (alpha_samples + res1)', lambda: ops.conv_nhwc_train(x192, 192, w_fc2, bias(96), out=o96, res1=y96, alpha1_samples=alpha), 192 + 192)
timed('proj fwd (alpha + res1 + res2*cs)', lambda: ops.conv_nhwc_train(x96, 96, w_proj, bias(96), out=o96, res1=y96, alpha1_samples=alpha, res2=z96, res2_cs=cs), 96 * 4)
timed('fc2 dgrad (gelu_grad_of)', lambda: ops.conv_nhwc_train(x96, 96, wd(96, 192), None, out=o192, gelu_grad_of=p192), 96 + 384)
timed('fc1 dgrad (plain)', lambda: ops.conv_nhwc(x192, 192, wd(192, 96), None, out=o96), 192 + 96)
timed('qkv dgrad (plain)', lambda: ops.conv_nhwc(x288, 288, wd(288, 96), None, out=o96), 288 + 96)
timed('cab.0 fwd (3x3 96->32 GELU + pre_out)', lambda: ops.conv_nhwc_train(x96, 96, wc(32, 96), bias(32), out=o32, act=L.ACT_GELU, pre_out=pre32), 96 + 64)
timed('cab.2 dgrad (3x3 96->32 gelu_grad_of)', lambda: ops.conv_nhwc_train(x96, 96, wdc(96, 32), None, out=o32, gelu_grad_of=pre32), 96 + 64)
timed('cab.0 dgrad (3x3 32->96 += )', lambda: ops.conv_nhwc(x32, 32, wdc(32, 96), None, out=o96, res1=o96), 32 + 192)
for name, xx, ci, dd, co, taps in [('qkv', x96, 96, x288, 288, 1), ('proj', x96, 96, y96, 96, 1), ('fc1', x96, 96, x192, 192, 1),
                                   ('fc2', x192, 192, y96, 96, 1), ('cab.0', x96, 96, x32, 32, 9), ('cab.2', x32, 32, y96, 96, 9),
                                   ('conv 96->96', x96, 96, y96, 96, 9)]:
    timed(f'wgrad {name} ({ci}->{co}, {taps} tap)', lambda: ops.conv_wgrad(xx, ci, dd, co, taps), ci + co)
timed('layernorm96', lambda: ops.layernorm96(x96.view(-1, 96), alpha.new_ones(96), alpha.new_zeros(96)), 192)
timed('layernorm96_bwd (+res)', lambda: ops.layernorm96_bwd(x96.view(-1, 96), y96.view(-1, 96), alpha.new_ones(96), z96.view(-1, 96)), 384)
qkv = x288
# window attention of one HAB (all 16384 windows x 6 heads), forward and backward
table = torch.randn(225, 6, device=dev) * 0.5
coords = torch.stack(torch.meshgrid(torch.arange(8), torch.arange(8), indexing='ij')).flatten(1)
rel = (coords[:, :, None] - coords[:, None, :]).permute(1, 2, 0) + 7
index = (rel[..., 0] * 15 + rel[..., 1]).to(dev)
wbias = ops.gather_rel_bias(table, index)
timed('window attention fwd (shift 4)', lambda: ops.window_attention(qkv, wbias, 4), 288 + 96)
timed('window attention bwd (shift 4)', lambda: ops.window_attention_bwd(qkv, wbias, index, y96, 4), 288 + 96 + 288)
