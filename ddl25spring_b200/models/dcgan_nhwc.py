"""DCGAN 64x64 on the channels-last tcgen05 convolution stack (kernel-native twin of models/dcgan.py).

Generator:      z [N, nz] --Linear (= ConvTranspose 4x4 on a 1x1 map, tcgen05 GEMM)--> [N,4,4,8g] -> BN+ReLU
                -> 3 x (ConvTranspose 4x4/2 as a four-phase tap GEMM -> BN+ReLU) -> ConvTranspose 4x4/2 + tanh (fused epilogue)
Discriminator:  image -> conv 4x4/2 + LeakyReLU (fused epilogue) -> 3 x (conv 4x4/2 tap GEMM on the input's even/odd phases
                -> BN+LeakyReLU fused) -> Linear over the 4x4x8d map (= the final 4x4 valid convolution)

Module / parameter names match models/dcgan.py (``blocks.<i>.0.weight`` ...); weights are stored in the layouts the
kernels read: convolutions OHWI ``[Cout, kh, kw, Cin]``, transposed convolutions ``[Cin, kh, kw, Cout]`` (the OHWI weight
of the convolution they transpose), the generator's first layer ``[kh, kw, Cout, nz]``, image channels padded 3 -> 4.
``from_torch_layout`` converts a models/dcgan.py state dict.  Images enter and leave as NCHW [N, 3, 64, 64].
[north star: BASELINE.json config 5; SURVEY §2.7]
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from ..ops import conv as K
from ..ops.nn import kernel_linear


class _Weight(nn.Module):
    def __init__(self, *shape, std: float = 0.02, zero_tail: int = 0):
        super().__init__()
        w = torch.empty(*shape).normal_(0.0, std)
        if zero_tail:
            w[..., -zero_tail:] = 0
        self.weight = nn.Parameter(w)


def _bn(c: int) -> nn.BatchNorm2d:
    bn = nn.BatchNorm2d(c)
    nn.init.normal_(bn.weight, 1.0, 0.02)
    nn.init.zeros_(bn.bias)
    return bn


class GeneratorNHWC(nn.Module):
    def __init__(self, nz: int = 100, ngf: int = 64, nc: int = 3):
        super().__init__()
        assert nz % 4 == 0 and nc <= 4
        self.nz, self.nc = nz, nc
        chans = [ngf * 8, ngf * 4, ngf * 2, ngf]
        blocks = [nn.Sequential(_Weight(4, 4, chans[0], nz), _bn(chans[0]))]
        blocks += [nn.Sequential(_Weight(a, 4, 4, b), _bn(b)) for a, b in zip(chans[:-1], chans[1:])]
        blocks.append(nn.Sequential(_Weight(ngf, 4, 4, 4, zero_tail=4 - nc)))
        self.blocks = nn.Sequential(*blocks)

    def _block(self, i: int, y):
        blk = self.blocks[i]
        if i == 0:
            z = y.reshape(y.shape[0], self.nz)
            w0 = blk[0].weight
            y = kernel_linear(z, w0.view(-1, self.nz), None, precise=False).view(z.shape[0], 4, 4, w0.shape[2])
            return K.batch_norm_nhwc(y, blk[1], relu=True)
        if i == len(self.blocks) - 1:
            img = K.conv_transpose2d_nhwc(y, blk[0].weight, None, 2, 1, act="tanh")
            return img[..., :self.nc].permute(0, 3, 1, 2)
        return K.batch_norm_nhwc(K.conv_transpose2d_nhwc(y, blk[0].weight, None, 2, 1), blk[1], relu=True)

    def forward(self, z):
        y = z
        for i in range(len(self.blocks)):
            y = self._block(i, y)
        return y

    def stages(self, cut: int = 3):
        """Two pipeline stages (blocks [0, cut) | [cut, end)); NHWC activations travel between them."""
        return _Stage(self, range(0, cut)), _Stage(self, range(cut, len(self.blocks)))


class DiscriminatorNHWC(nn.Module):
    def __init__(self, ndf: int = 64, nc: int = 3):
        super().__init__()
        assert nc <= 4
        self.nc = nc
        chans = [ndf, ndf * 2, ndf * 4, ndf * 8]
        blocks = [nn.Sequential(_Weight(ndf, 4, 4, 4, zero_tail=4 - nc))]
        blocks += [nn.Sequential(_Weight(b, 4, 4, a), _bn(b)) for a, b in zip(chans[:-1], chans[1:])]
        blocks.append(nn.Sequential(_Weight(1, 4, 4, chans[-1])))
        self.blocks = nn.Sequential(*blocks)

    def _block(self, i: int, y):
        blk = self.blocks[i]
        if i == 0:
            x = F.pad(y.permute(0, 2, 3, 1), (0, 4 - self.nc))                   # NCHW image -> NHWC, 4 channels (16-byte pixels)
            return K.conv2d_nhwc(x, blk[0].weight, 2, 1, act="leaky_relu", slope=0.2)
        if i == len(self.blocks) - 1:
            w = blk[0].weight
            return kernel_linear(y.reshape(y.shape[0], -1), w.view(1, -1), None, precise=False).squeeze(1)   # logits
        return K.batch_norm_nhwc(K.conv2d_nhwc(y, blk[0].weight, 2, 1), blk[1], leaky=0.2)

    def forward(self, x):
        y = x
        for i in range(len(self.blocks)):
            y = self._block(i, y)
        return y

    def stages(self, cut: int = 2):
        return _Stage(self, range(0, cut)), _Stage(self, range(cut, len(self.blocks)))


class _Stage(nn.Module):
    """A contiguous run of blocks of a GeneratorNHWC / DiscriminatorNHWC as its own module (pipeline stage)."""

    def __init__(self, net: nn.Module, idx):
        super().__init__()
        self.idx = list(idx)
        self.blocks = nn.ModuleList(net.blocks[i] for i in self.idx)      # registers exactly this stage's parameters
        object.__setattr__(self, "_net", net)                             # not a submodule: no parameter duplication

    def forward(self, y):
        for i in self.idx:
            y = self._net._block(i, y)
        return y


def from_torch_layout(sd: dict, kind: str) -> dict:
    """State dict of models/dcgan.py ``Generator`` (kind='G') or ``Discriminator`` (kind='D') -> the NHWC twin's layout."""
    out = {}
    last = max(int(k.split(".")[1]) for k in sd)
    for k, v in sd.items():
        i = int(k.split(".")[1])
        if v.dim() == 4:
            if kind == "G":
                # ConvTranspose2d weight [Cin, Cout, kh, kw]
                if i == 0:
                    v = v.permute(2, 3, 1, 0).contiguous()                           # [kh, kw, Cout, nz]
                else:
                    v = v.permute(0, 2, 3, 1).contiguous()                           # [Cin, kh, kw, Cout]
                    if i == last and v.shape[-1] < 4:
                        v = F.pad(v, (0, 4 - v.shape[-1]))
            else:
                v = v.permute(0, 2, 3, 1).contiguous()                               # Conv2d [Cout, Cin, kh, kw] -> OHWI
                if i == 0 and v.shape[-1] < 4:
                    v = F.pad(v, (0, 4 - v.shape[-1]))
        out[k] = v
    return out
