"""ResNet-18 on the channels-last tcgen05 convolution stack (kernel-native twin of models/resnet.py).

Same network, same parameter *names* as ``models.resnet.ResNet18`` / torchvision's ``resnet18`` (so arenas, checkpoints
and aggregation rules treat both alike), but convolution weights are stored OHWI ``[Cout, kh, kw, Cin]`` — the layout
the tap-GEMM kernels read with TMA — and activations stay NHWC from the stem to the classifier:

  stem     im2col(7x7/2) + tcgen05 GEMM           -> BN+ReLU (fused) -> 3x3/2 max pool
  blocks   conv3x3 (tap GEMM) -> BN+ReLU -> conv3x3 -> BN + shortcut + ReLU (one fused pass)
           stride-2 blocks read the even/odd phases of their input; the 1x1/2 shortcut conv is a one-tap GEMM
  head     global average pool -> tcgen05 Linear

``from_torch_layout`` / ``to_torch_layout`` convert state dicts to and from the NCHW-weight model.  On CPU the ops fall
back to plain PyTorch, so both models agree to rounding there (tests/test_conv_nhwc.py).
[north star: BASELINE.json configs 2 and 4; SURVEY §2.7]
"""
from __future__ import annotations

import math

import torch
from torch import nn

from ..ops import conv as K
from ..ops.nn import kernel_linear


class ConvNHWC(nn.Module):
    """Bias-free convolution with an OHWI weight."""

    def __init__(self, cin: int, cout: int, k: int, stride: int, padding: int):
        super().__init__()
        self.stride, self.padding = stride, padding
        self.weight = nn.Parameter(torch.empty(cout, k, k, cin))
        nn.init.normal_(self.weight, 0.0, math.sqrt(2.0 / (cout * k * k)))           # kaiming_normal_, fan_out, relu

    def forward(self, x, phases=None):
        return K.conv2d_nhwc(x, self.weight, self.stride, self.padding, phases)


class StemConv(nn.Module):
    """7x7 stride-2 convolution of the NCHW input image; the weight carries a zero 4th input channel (16-byte taps)."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        assert cin <= 4
        self.cin = cin
        w = torch.zeros(cout, 7, 7, 4)
        w[..., :cin].normal_(0.0, math.sqrt(2.0 / (cout * 49)))
        self.weight = nn.Parameter(w)

    def forward(self, x):
        return K.stem_conv7x7s2(x, self.weight)


class BasicBlockNHWC(nn.Module):
    expansion = 1

    def __init__(self, cin: int, cout: int, stride: int = 1):
        super().__init__()
        self.conv1 = ConvNHWC(cin, cout, 3, stride, 1)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = ConvNHWC(cout, cout, 3, 1, 1)
        self.bn2 = nn.BatchNorm2d(cout)
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(ConvNHWC(cin, cout, 1, stride, 0), nn.BatchNorm2d(cout))

    def forward(self, x):
        idt = x
        ph = K.stride2_phases(x) if self.conv1.stride == 2 else None      # one phase split feeds conv1 and the shortcut conv
        if self.downsample is not None:
            idt = K.batch_norm_nhwc(self.downsample[0](x, ph if self.downsample[0].stride == 2 else None), self.downsample[1])
        y = K.batch_norm_nhwc(self.conv1(x, ph), self.bn1, relu=True)
        return K.batch_norm_nhwc(self.conv2(y), self.bn2, relu=True, residual=idt)


class ResNet18NHWC(nn.Module):
    def __init__(self, num_classes: int = 10, in_ch: int = 3):
        super().__init__()
        self.conv1 = StemConv(in_ch, 64)
        self.bn1 = nn.BatchNorm2d(64)
        widths, cin = (64, 128, 256, 512), 64
        for i, w in enumerate(widths, start=1):
            stride = 1 if i == 1 else 2
            setattr(self, f"layer{i}", nn.Sequential(BasicBlockNHWC(cin, w, stride), BasicBlockNHWC(w, w, 1)))
            cin = w
        self.fc = nn.Linear(512, num_classes)

    def forward(self, x):
        """``x``: NCHW image batch [N, 3, H, W] (H, W multiples of 32) -> logits [N, num_classes]."""
        y = K.batch_norm_nhwc(self.conv1(x), self.bn1, relu=True)
        y = K.max_pool3x3s2_nhwc(y)
        for i in range(1, 5):
            y = getattr(self, f"layer{i}")(y)
        return kernel_linear(K.global_avg_pool_nhwc(y), self.fc.weight, self.fc.bias, precise=False)


def resnet18_nhwc(num_classes: int = 10) -> ResNet18NHWC:
    return ResNet18NHWC(num_classes)


def split_resnet18_nhwc_two_stage(model: ResNet18NHWC) -> tuple[nn.Module, nn.Module]:
    """2-stage model-parallel split (BASELINE.json config 5): stem+layer1+layer2 | layer3+layer4+fc; the NHWC activation
    [N, H/8, W/8, 128] travels between the stages (same cut as models.resnet.split_resnet18_two_stage)."""

    class Front(nn.Module):
        def __init__(s):
            super().__init__()
            s.conv1, s.bn1, s.layer1, s.layer2 = model.conv1, model.bn1, model.layer1, model.layer2

        def forward(s, x):
            y = K.max_pool3x3s2_nhwc(K.batch_norm_nhwc(s.conv1(x), s.bn1, relu=True))
            return s.layer2(s.layer1(y))

    class Back(nn.Module):
        def __init__(s):
            super().__init__()
            s.layer3, s.layer4, s.fc = model.layer3, model.layer4, model.fc

        def forward(s, y):
            y = s.layer4(s.layer3(y))
            return kernel_linear(K.global_avg_pool_nhwc(y), s.fc.weight, s.fc.bias, precise=False)

    return Front(), Back()


def from_torch_layout(sd: dict) -> dict:
    """State dict of ``models.resnet.ResNet18`` (OIHW conv weights) -> ``ResNet18NHWC`` (OHWI; stem padded to 4 channels)."""
    out = {}
    for k, v in sd.items():
        if v.dim() == 4:
            v = v.permute(0, 2, 3, 1).contiguous()
            if k == "conv1.weight" and v.shape[-1] < 4:
                v = torch.nn.functional.pad(v, (0, 4 - v.shape[-1]))
        out[k] = v
    return out


def to_torch_layout(sd: dict, in_ch: int = 3) -> dict:
    out = {}
    for k, v in sd.items():
        if v.dim() == 4:
            if k == "conv1.weight":
                v = v[..., :in_ch]
            v = v.permute(0, 3, 1, 2).contiguous()
        out[k] = v
    return out
