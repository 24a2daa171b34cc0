"""Oracle arithmetic with the kernels' rounding points, shared by the gradient-parity tests (test infrastructure)."""
import torch


def _q(t):
    """bf16 rounding with a straight-through gradient (the kernels store activations in bf16)"""
    return t + (t.bfloat16().float() - t).detach()


def _emulated_rrdbnet(net, x):
    """oracle RRDBNet arithmetic with the kernels' rounding points: bf16 conv operands (activations and weights),
    fp32 accumulation, bf16-stored activations, fp32 residual trunk.  Same LeakyReLU sign pattern as the libsrb forward,
    so the comparison isolates the backward kernels from forward rounding (sign flips at random init)."""
    import torch.nn.functional as F
    conv = lambda m, t: F.conv2d(t, _q(m.weight), m.bias, padding=1)
    lr = lambda t: F.leaky_relu(t, 0.2)
    t32 = conv(net.conv_first, _q(F.pixel_unshuffle(x, 2)))
    f0 = t32
    for rrdb in net.body:
        t_rrdb = t32
        for rdb in (rrdb.rdb1, rrdb.rdb2, rrdb.rdb3):
            xs = [_q(t32)]
            for c in (rdb.conv1, rdb.conv2, rdb.conv3, rdb.conv4):
                xs.append(_q(lr(conv(c, torch.cat(xs, 1)))))
            t32 = conv(rdb.conv5, torch.cat(xs, 1)) * 0.2 + t32
        t32 = t32 * 0.2 + t_rrdb
    feat = _q(conv(net.conv_body, _q(t32)) + f0)
    up = lambda t: F.interpolate(t, scale_factor=2, mode='nearest')
    a1 = _q(lr(conv(net.conv_up1, up(feat))))
    a2 = _q(lr(conv(net.conv_up2, up(a1))))
    a3 = _q(lr(conv(net.conv_hr, a2)))
    return F.conv2d(a3, net.conv_last.weight, net.conv_last.bias, padding=1)
