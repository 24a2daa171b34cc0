"""Seeded 'trained-like' parameters for the parity tests: at the library's random init the model
is nearly linear (RDB convs x0.1, every bias 0, LayerNorm gamma=1 / beta=0, relative-position tables ~0.02) and the
final output has sigma ~0.015, so a PSNR bar on the [0,1] scale cannot see a wrong bias table, a wrong shift mask or a
dead CAB branch.  This perturbation switches every one of those paths on while keeping activations O(1):
  RDB conv weights x rdb_gain, every conv / linear bias ~ N(0, bias_std), LayerNorm gamma = 1 +- 0.2, beta +- 0.1,
  both relative-position tables ~ N(0, table_std), channel-attention 1x1 convs and the CAB 3x3 convs amplified (the CAB
  enters the stream x 0.01), HAT linears x lin_gain, stage2.conv_last x out_gain (final sigma ~0.1).
"""
import torch


def make_trained_like(model, seed=1234, rdb_gain=6.0, bias_std=0.1, lin_gain=6.0, table_std=1.5, cab_gain=6.0,
                      out_gain=8.0):
    g = torch.Generator().manual_seed(seed)

    def rn(t, std):
        return torch.randn(t.shape, generator=g) * std

    with torch.no_grad():
        for name, p in model.named_parameters():
            if name.startswith('stage1.body.') and name.endswith('.weight'):
                p.mul_(rdb_gain)
            elif 'relative_position_bias_table' in name:
                p.copy_(rn(p, table_std * (2.5 if 'overlap_attn' in name else 1.0)))
            elif '.norm' in name and name.endswith('.weight'):
                p.copy_(1.0 + rn(p, 0.2))
            elif '.norm' in name and name.endswith('.bias'):
                p.copy_(rn(p, 0.1))
            elif 'cab.3.attention' in name and name.endswith('.weight'):
                p.copy_(rn(p, 1.0))
            elif ('cab.0.weight' in name) or ('cab.2.weight' in name):  # CAB enters the stream x conv_scale = 0.01
                p.mul_(cab_gain)
            elif name == 'stage2.conv_last.weight':  # final signal sigma ~0.1 instead of ~0.01
                p.mul_(out_gain)
            elif name.endswith('.bias'):
                p.copy_(rn(p, bias_std))
            elif name.startswith('stage2.') and p.dim() == 2:  # HAT linears (trunc-normal 0.02 at init)
                p.mul_(lin_gain)
    return model
