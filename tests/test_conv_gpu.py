"""Parity of the tcgen05 implicit-GEMM conv (C ABI srb_conv_nhwc_bf16) against F.conv2d / F.linear in fp32 on the
same bf16-rounded inputs and weights.  Tolerance: fp32 accumulation -> differences come from accumulation order and
the final bf16 rounding of the output only (rel 2^-8)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref(x_nhwc, w, bias, taps):
    x = x_nhwc.float().permute(0, 3, 1, 2)
    wq = w.to(torch.bfloat16).float()
    if taps == 9:
        y = F.conv2d(x, wq, bias, padding=1)
    else:
        y = F.conv2d(x, wq.view(w.shape[0], w.shape[1], 1, 1), bias)
    return y.permute(0, 2, 3, 1).contiguous()


def _close(got, ref, what):
    err = (got.float() - ref).abs()
    tol = 1e-2 * ref.abs() + 2e-2 * ref.abs().mean() + 1e-3
    bad = (err > tol).sum().item()
    assert bad == 0, f'{what}: {bad} mismatches, max err {err.max().item():.4g}, ref absmax {ref.abs().max().item():.4g}'


CASES = [
    # B, H, W, Cin, Cout, taps, in_pitch
    (2, 64, 64, 64, 32, 9, 192),
    (2, 64, 64, 96, 32, 9, 192),
    (1, 32, 64, 128, 32, 9, 192),
    (1, 32, 32, 160, 32, 9, 192),
    (2, 64, 64, 192, 64, 9, 192),
    (3, 16, 32, 64, 64, 9, 64),
    (1, 32, 32, 96, 96, 9, 96),
    (1, 48, 80, 32, 96, 9, 32),
    (1, 32, 32, 96, 64, 9, 96),
    (1, 32, 32, 96, 288, 1, 96),
    (1, 32, 32, 192, 96, 1, 192),
    (1, 16, 16, 96, 192, 1, 96),
]


@pytest.mark.parametrize('B,H,W,Cin,Cout,taps,pitch', CASES)
def test_conv_plain(cuda, B, H, W, Cin, Cout, taps, pitch):
    from superresolution_b200 import ops, _lib as L
    g = torch.Generator(device='cpu').manual_seed(Cin * 1000 + Cout + taps)
    buf = torch.randn(B, H, W, pitch, generator=g).to(cuda).to(torch.bfloat16)
    k = 3 if taps == 9 else 1
    w = (torch.randn(Cout, Cin, k, k, generator=g) / (Cin * taps) ** 0.5).to(cuda)
    bias = torch.randn(Cout, generator=g).to(cuda)
    wp = ops.pack_conv_weight(w)
    out = torch.full((B, H, W, Cout), float('nan'), device=cuda, dtype=torch.bfloat16)
    ops.conv_nhwc(buf, Cin, wp, bias, out=out)
    torch.cuda.synchronize()
    ref = _ref(buf[..., :Cin], w, bias, taps)
    _close(out, ref, 'plain')


def test_conv_rdb_epilogue(cuda):
    """conv5 of an RDB: reads 192 channels, *0.2 + fp32 trunk, *0.2 + fp32 RRDB input, writes fp32 trunk and the
    bf16 slice [0,64) of the next dense-block buffer."""
    from superresolution_b200 import ops, _lib as L
    g = torch.Generator(device='cpu').manual_seed(5)
    B, H, W = 2, 64, 64
    buf = torch.randn(B, H, W, 192, generator=g).to(cuda).to(torch.bfloat16)
    w = (torch.randn(64, 192, 3, 3, generator=g) / 40).to(cuda)
    bias = torch.randn(64, generator=g).to(cuda)
    r1 = torch.randn(B, H, W, 64, generator=g).to(cuda)
    r2 = torch.randn(B, H, W, 64, generator=g).to(cuda)
    nxt = torch.zeros(B, H, W, 192, device=cuda, dtype=torch.bfloat16)
    o32 = torch.zeros(B, H, W, 64, device=cuda)
    ops.conv_nhwc(buf, 192, ops.pack_conv_weight(w), bias, out=nxt[..., :64], out32=o32, alpha1=0.2, res1=r1,
                  alpha2=0.2, res2=r2)
    torch.cuda.synchronize()
    ref = (_ref(buf, w, bias, 9) * 0.2 + r1) * 0.2 + r2
    _close(o32, ref, 'fp32 trunk')
    _close(nxt[..., :64], ref, 'bf16 slice')
    assert nxt[..., 64:].abs().max().item() == 0


def test_conv_lrelu_slice(cuda):
    """conv2 of an RDB: reads channels [0,96) of the 192-pitch buffer, LeakyReLU(0.2), writes slice [96,128)."""
    from superresolution_b200 import ops, _lib as L
    g = torch.Generator(device='cpu').manual_seed(6)
    B, H, W = 1, 64, 64
    buf = torch.randn(B, H, W, 192, generator=g).to(cuda).to(torch.bfloat16)
    before = buf.clone()
    w = (torch.randn(32, 96, 3, 3, generator=g) / 30).to(cuda)
    bias = torch.randn(32, generator=g).to(cuda)
    ops.conv_nhwc(buf, 96, ops.pack_conv_weight(w), bias, out=buf[..., 96:128], act=L.ACT_LRELU, slope=0.2)
    torch.cuda.synchronize()
    ref = F.leaky_relu(_ref(before[..., :96], w, bias, 9), 0.2)
    _close(buf[..., 96:128], ref, 'slice')
    assert torch.equal(buf[..., :96], before[..., :96]) and torch.equal(buf[..., 128:], before[..., 128:])


def test_conv_gelu_and_scaled_residual(cuda):
    """HAB-style epilogues: GELU (fc1) and proj + shortcut + cab * per-(image,channel) scale."""
    from superresolution_b200 import ops, _lib as L
    g = torch.Generator(device='cpu').manual_seed(7)
    B, H, W = 2, 32, 32
    x = torch.randn(B, H, W, 96, generator=g).to(cuda).to(torch.bfloat16)
    w1 = (torch.randn(192, 96, generator=g) / 10).to(cuda)
    b1 = torch.randn(192, generator=g).to(cuda)
    h = torch.empty(B, H, W, 192, device=cuda, dtype=torch.bfloat16)
    ops.conv_nhwc(x, 96, ops.pack_conv_weight(w1), b1, out=h, act=L.ACT_GELU)
    _close(h, F.gelu(_ref(x, w1, b1, 1)), 'gelu')
    wp = (torch.randn(96, 96, generator=g) / 10).to(cuda)
    bp = torch.randn(96, generator=g).to(cuda)
    sc = torch.randn(B, H, W, 96, generator=g).to(cuda).to(torch.bfloat16)
    cab = torch.randn(B, H, W, 96, generator=g).to(cuda).to(torch.bfloat16)
    cs = torch.rand(B, 96, generator=g).to(cuda)
    out = torch.empty(B, H, W, 96, device=cuda, dtype=torch.bfloat16)
    ops.conv_nhwc(x, 96, ops.pack_conv_weight(wp), bp, out=out, res1=sc, res2=cab, res2_cs=cs)
    ref = _ref(x, wp, bp, 1) + sc.float() + cab.float() * cs[:, None, None, :]
    _close(out, ref, 'combine')


def test_conv_pixel_shuffle(cuda):
    from superresolution_b200 import ops
    g = torch.Generator(device='cpu').manual_seed(8)
    B, H, W = 1, 32, 32
    x = torch.randn(B, H, W, 64, generator=g).to(cuda).to(torch.bfloat16)
    w = (torch.randn(256, 64, 3, 3, generator=g) / 24).to(cuda)
    bias = torch.randn(256, generator=g).to(cuda)
    out = torch.empty(B, 2 * H, 2 * W, 64, device=cuda, dtype=torch.bfloat16)
    ops.conv_nhwc(x, 64, ops.pack_conv_weight(w), bias, out=out, pixel_shuffle=True)
    ref = F.pixel_shuffle(_ref(x, w, bias, 9).permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    _close(out, ref, 'pixel shuffle')
