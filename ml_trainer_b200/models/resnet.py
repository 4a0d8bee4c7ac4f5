"""ResNet-50 (BASELINE.json config 3) on the op library.

Module tree, parameter / buffer names and shapes follow ``torchvision.models.resnet50`` (v1.5: the stride sits on
the 3x3 convolution; 25,557,032 parameters, 53 BatchNorm layers), so state dicts are interchangeable and torchvision
is the numerical oracle in the tests.  Input is the usual NCHW float batch.

GPU path: activations are NHWC bf16; every convolution is (im2col +) a tcgen05 GEMM with its weight held physically
channels-last in the optimizer arena; BatchNorm runs as stats + apply kernels with ReLU and the residual add fused into
the apply pass; 3x3/2 max-pool, global average pool and the 2048->1000 classifier GEMM complete the network.
CPU path: the same modules through plain ``torch.nn.functional`` calls.
"""
from __future__ import annotations

import torch
import torch.nn.functional as TF
from torch import nn

from ..ops import functional as F
from . import register


def _conv(cin, cout, k, stride=1, pad=0):
    m = nn.Conv2d(cin, cout, k, stride=stride, padding=pad, bias=False)
    m.weight._mlt_channels_last = True      # FusedOptimizer lays it out as [Cout, KH, KW, Cin]
    return m


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _conv(inplanes, planes, 1)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = _conv(planes, planes, 3, stride, 1)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = _conv(planes, planes * 4, 1)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):                       # CPU / oracle path (NCHW)
        idt = x
        o = self.relu(self.bn1(self.conv1(x)))
        o = self.relu(self.bn2(self.conv2(o)))
        o = self.bn3(self.conv3(o))
        if self.downsample is not None:
            idt = self.downsample(x)
        return self.relu(o + idt)

    def forward_nhwc(self, x):
        from ..ops import conv as C

        # identity skip: the block input gets its skip gradient added inside conv1's data-gradient GEMM (ops/conv.py::SkipGrad)
        skip = None
        if (self.downsample is None and self.bn3.training and torch.is_grad_enabled() and x.requires_grad
                and F.SKIP_GRAD_FUSION):
            skip = C.SkipGrad()
        o = C.batch_norm_nhwc(C.conv2d_nhwc(x, self.conv1.weight, skip=skip), self.bn1, relu=True)
        o = C.batch_norm_nhwc(C.conv2d_nhwc(o, self.conv2.weight, self.stride, 1), self.bn2, relu=True)
        o = C.conv2d_nhwc(o, self.conv3.weight)
        if self.downsample is not None:
            idt = C.batch_norm_nhwc(C.conv2d_nhwc(x, self.downsample[0].weight, self.stride, 0), self.downsample[1])
        else:
            idt = x
        return C.batch_norm_nhwc(o, self.bn3, relu=True, residual=idt, skip=skip)      # relu(bn3(conv3) + identity), one pass


class ResNet(nn.Module):
    _mlt_native = True
    _mlt_accepts_fp32_input = True        # the NHWC conversion kernel reads the fp32 NCHW batch directly

    def __init__(self, layers=(3, 4, 6, 3), num_classes=1000, width=64):
        super().__init__()
        self.inplanes = width
        self.conv1 = _conv(3, width, 7, 2, 3)
        self.bn1 = nn.BatchNorm2d(width)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(width, layers[0])
        self.layer2 = self._make_layer(width * 2, layers[1], 2)
        self.layer3 = self._make_layer(width * 4, layers[2], 2)
        self.layer4 = self._make_layer(width * 8, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(width * 8 * 4, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def _make_layer(self, planes, blocks, stride=1):
        down = None
        if stride != 1 or self.inplanes != planes * 4:
            down = nn.Sequential(_conv(self.inplanes, planes * 4, 1, stride), nn.BatchNorm2d(planes * 4))
        layers = [Bottleneck(self.inplanes, planes, stride, down)]
        self.inplanes = planes * 4
        layers += [Bottleneck(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x):
        if x.is_cuda and F.native_available() and self.conv1.weight.dtype == torch.bfloat16:
            return self._forward_native(x)
        return self._forward_torch(x.to(self.conv1.weight.dtype))

    def _forward_torch(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x)

    def _bump_batch_counters(self):
        """``num_batches_tracked += 1`` for all 53 BatchNorm layers as one multi-tensor call instead of 53 one-element kernels; the
        layers are told that their counter has been taken care of for this forward pass."""
        bns = self.__dict__.get("_bn_layers")
        if bns is None:
            bns = [m for m in self.modules() if isinstance(m, nn.BatchNorm2d) and m.num_batches_tracked is not None]
            self.__dict__["_bn_layers"] = bns             # (plain attribute: not a sub-module registration)
        live = [m for m in bns if m.training]
        if live:
            torch._foreach_add_([m.num_batches_tracked for m in live], 1)
        for m in bns:
            m._mlt_counter_done = m.training

    def _forward_native(self, x):
        from ..ops import conv as C

        self._bump_batch_counters()
        h = C.to_nhwc_bf16(x, 8)                                            # [B, H, W, 8], channels 3..7 are zero
        w = C.pad_weight_channels(self.conv1.weight, 8)                      # stem filter padded the same way
        h = C.batch_norm_nhwc(C.conv2d_nhwc(h, w, 2, 3), self.bn1, relu=True)
        h = C.max_pool_nhwc(h, 3, 2, 1)
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
            for blk in layer:
                h = blk.forward_nhwc(h)
        h = C.global_avg_pool_nhwc(h)
        return F.linear_act(h, self.fc.weight, self.fc.bias).float()


@register("resnet50")
def resnet50(num_classes: int = 1000, **kw):
    return ResNet((3, 4, 6, 3), num_classes=num_classes, **kw)


@register("resnet_tiny")
def resnet_tiny(num_classes: int = 16, **kw):
    """One bottleneck per stage, width 16: the same code paths at test size."""
    return ResNet((1, 1, 1, 1), num_classes=num_classes, width=kw.pop("width", 16), **kw)
