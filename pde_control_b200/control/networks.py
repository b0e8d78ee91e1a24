"""Control networks as ordinary PyTorch modules (the physics is the product; the nets are not re-engineered).

Architectures restated from the reference (weights are NOT compatible with the TF checkpoints):
  OPResNet  1-D observation predictor  src/control/pde/burgers.py:45-80 (op_resnet) with residual_block_1d
            PF/tf/util.py:148-180
  UNet      2-D u_net                  PF/tf/standard_networks.py:9-69 with conv_layer / residual_block
            PF/tf/grid_layers.py:6-51 and math.upsample2x PF/math/nd.py:353-368
TF 'SYMMETRIC' padding of width 1 == torch 'replicate'; upsample2x (0.25/0.75 weights on a replicate-padded tensor) ==
F.interpolate(scale_factor=2, mode='(bi)linear', align_corners=False); the 1-D case is written out (upsample2x_1d).
Tensors enter and leave in the reference's channels-last layout.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


def upsample2x_1d(y):
    """math.upsample2x on [B, C, N] (PF/math/nd.py:353-368): 0.25/0.75 weights on a replicate-padded tensor.

    Same values as F.interpolate(scale_factor=2, mode='linear', align_corners=False), written with slices because
    aten's upsample_linear1d kernels serialise over batch x channels (12 ms per call at [4096, 16, 80] on a B200,
    80 % of an OPResNet forward+backward; profiles/r1j_opnet_profile.md).
    """
    p = F.pad(y, (1, 1), mode="replicate")
    even = 0.25 * p[..., :-2] + 0.75 * y
    odd = 0.75 * y + 0.25 * p[..., 2:]
    return torch.stack([even, odd], -1).flatten(-2)


class _ResBlock(nn.Module):
    def __init__(self, dims, cin, cout):
        nn.Module.__init__(self)
        conv = nn.Conv1d if dims == 1 else nn.Conv2d
        self.conv1 = conv(cin, cout, 3)
        self.conv2 = conv(cout, cout, 3)
        self.pad = (1, 1) if dims == 1 else (1, 1, 1, 1)

    def forward(self, y):
        s = y
        y = F.leaky_relu(self.conv1(F.pad(y, self.pad, mode="replicate")), 0.2)     # tf.nn.leaky_relu default alpha 0.2
        y = self.conv2(F.pad(y, self.pad, mode="replicate"))
        return F.leaky_relu(y + s, 0.2)


class OPResNet(nn.Module):
    """Predicts the centre frame from (initial, target): [B, N, 1] x 2 -> [B, N, 1]."""

    FILTERS = (4, 8, 16, 16, 16)

    def __init__(self):
        nn.Module.__init__(self)
        self.down = nn.ModuleList()
        self.down_rb = nn.ModuleList()
        cin = 2
        for f in self.FILTERS:
            self.down.append(nn.Conv1d(cin, f, 2, stride=2))
            self.down_rb.append(nn.Sequential(_ResBlock(1, f, f), _ResBlock(1, f, f)))
            cin = f
        self.center = nn.Sequential(*[_ResBlock(1, 16, 16) for _ in range(3)])
        skip_channels = [16, 16, 8, 4, 2]                    # resolutions[1:]: the four coarser down outputs, then the input
        self.up = nn.ModuleList()
        self.up_rb = nn.ModuleList()
        for i, sc in enumerate(skip_channels):
            last = i == len(skip_channels) - 1
            self.up.append(nn.Conv1d(16 + sc, 1 if last else 16, 2))
            self.up_rb.append(nn.Identity() if last else nn.Sequential(_ResBlock(1, 16, 16), _ResBlock(1, 16, 16)))

    def forward(self, initial, target):
        y = torch.cat([initial, target], -1).transpose(1, 2)            # [B, 2, N]
        y = F.pad(y, (0, 31))                                           # downres_padding = 1+2+4+8+16
        res = [y]
        for conv, rb in zip(self.down, self.down_rb):
            y = rb(F.relu(conv(res[0])))
            res.insert(0, y)
        y = self.center(y)
        for i, skip in enumerate(res[1:]):
            y = upsample2x_1d(y)
            y = torch.cat([y, skip[:, :, :y.shape[2]]], 1)
            y = self.up[i](F.pad(y, (0, 1), mode="replicate"))
            if i < len(res) - 2:
                y = self.up_rb[i](F.relu(y))
        return y.transpose(1, 2)


class UNet(nn.Module):
    """u_net(domain, input_fields, output_field, levels, filters=16): [B, ny, nx, Cin] -> [B, ny, nx, Cout]."""

    def __init__(self, in_channels, out_channels=1, levels=2, filters=16, blocks_per_level=2):
        nn.Module.__init__(self)
        self.levels = levels
        self.down = nn.ModuleList()
        self.down_rb = nn.ModuleList()
        cin = in_channels
        skip = [in_channels]
        for _ in range(levels):
            self.down.append(nn.Conv2d(cin, filters, 2, stride=2))
            self.down_rb.append(nn.Sequential(*[_ResBlock(2, filters, filters) for _ in range(blocks_per_level)]))
            cin = filters
            skip.insert(0, filters)
        self.up = nn.ModuleList()
        self.up_rb = nn.ModuleList()
        for level in range(levels):
            last = level == levels - 1
            sc = skip[level + 1]
            self.up.append(nn.Conv2d(filters + sc, out_channels if last else filters, 2))
            self.up_rb.append(nn.Identity() if last else nn.Sequential(*[_ResBlock(2, filters, filters) for _ in range(blocks_per_level)]))

    def forward(self, x):
        ny, nx = x.shape[1:3]
        y = x.permute(0, 3, 1, 2)
        pad = sum(2 ** i for i in range(self.levels))
        y = F.pad(y, (0, pad, 0, pad))                 # the CFE/OP inputs have 'constant' extrapolation on OPEN domains
        res = [y]
        for conv, rb in zip(self.down, self.down_rb):
            y = rb(F.relu(conv(res[0])))
            res.insert(0, y)
        y = res.pop(0)
        for level in range(self.levels):
            y = F.interpolate(y, scale_factor=2, mode="bilinear", align_corners=False)
            skip = res.pop(0)
            y = torch.cat([y, skip[:, :, :y.shape[2], :y.shape[3]]], 1)
            y = self.up[level](F.pad(y, (0, 1, 0, 1)))
            if level < self.levels - 1:
                y = self.up_rb[level](F.relu(y))
        return y[:, :, :ny, :nx].permute(0, 2, 3, 1)
