"""DCGAN 64x64 generator / discriminator [north star: BASELINE.json config 5].

Standard shapes (Radford et al.): G: z[nz=100,1,1] -> 512x4x4 -> 256x8x8 -> 128x16x16 ->
64x32x32 -> 3x64x64 (tanh); D mirrors it with stride-2 4x4 convs and LeakyReLU(0.2).
Both are splittable into two pipeline stages for the model-parallel config.
"""
from __future__ import annotations

from torch import nn


def _g_block(cin, cout, k, s, p, last=False):
    layers = [nn.ConvTranspose2d(cin, cout, k, s, p, bias=False)]
    layers += [nn.Tanh()] if last else [nn.BatchNorm2d(cout), nn.ReLU(True)]
    return nn.Sequential(*layers)


def _d_block(cin, cout, k, s, p, bn=True, last=False):
    layers = [nn.Conv2d(cin, cout, k, s, p, bias=False)]
    if last:
        return nn.Sequential(*layers)
    if bn:
        layers.append(nn.BatchNorm2d(cout))
    layers.append(nn.LeakyReLU(0.2, inplace=True))
    return nn.Sequential(*layers)


class Generator(nn.Module):
    def __init__(self, nz: int = 100, ngf: int = 64, nc: int = 3):
        super().__init__()
        self.nz = nz
        chans = [nz, ngf * 8, ngf * 4, ngf * 2, ngf]
        blocks = [_g_block(chans[0], chans[1], 4, 1, 0)]
        blocks += [_g_block(a, b, 4, 2, 1) for a, b in zip(chans[1:-1], chans[2:])]
        blocks.append(_g_block(ngf, nc, 4, 2, 1, last=True))
        self.blocks = nn.Sequential(*blocks)
        self.apply(_dcgan_init)

    def forward(self, z):
        if z.dim() == 2:
            z = z[:, :, None, None]
        return self.blocks(z)

    def stages(self, cut: int = 3):
        return nn.Sequential(*self.blocks[:cut]), nn.Sequential(*self.blocks[cut:])


class Discriminator(nn.Module):
    def __init__(self, ndf: int = 64, nc: int = 3):
        super().__init__()
        chans = [nc, ndf, ndf * 2, ndf * 4, ndf * 8]
        blocks = [_d_block(chans[0], chans[1], 4, 2, 1, bn=False)]
        blocks += [_d_block(a, b, 4, 2, 1) for a, b in zip(chans[1:-1], chans[2:])]
        blocks.append(_d_block(ndf * 8, 1, 4, 1, 0, last=True))
        self.blocks = nn.Sequential(*blocks)
        self.apply(_dcgan_init)

    def forward(self, x):
        return self.blocks(x).flatten(1).squeeze(1)   # logits; pair with BCE-with-logits

    def stages(self, cut: int = 2):
        return nn.Sequential(*self.blocks[:cut]), nn.Sequential(*self.blocks[cut:])


def _dcgan_init(m: nn.Module) -> None:
    if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
        nn.init.normal_(m.weight, 0.0, 0.02)
    elif isinstance(m, nn.BatchNorm2d):
        nn.init.normal_(m.weight, 1.0, 0.02)
        nn.init.zeros_(m.bias)
