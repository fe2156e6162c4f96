"""MinkowskiEngine.modules.resnet_block (A.7): BasicBlock / Bottleneck with ME's attribute names."""
import torch.nn as nn

from ... import ops
from .. import MinkowskiBatchNorm, MinkowskiConvolution, MinkowskiReLU


def _same_map(a, b):
    return a.coordinate_manager is b.coordinate_manager and a.coordinate_map_key == b.coordinate_map_key


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, dilation=1, downsample=None, bn_momentum=0.1, dimension=-1):
        super().__init__()
        assert dimension > 0
        self.conv1 = MinkowskiConvolution(inplanes, planes, kernel_size=3, stride=stride, dilation=dilation,
                                          dimension=dimension)
        self.norm1 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.conv2 = MinkowskiConvolution(planes, planes, kernel_size=3, stride=1, dilation=dilation,
                                          dimension=dimension)
        self.norm2 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.relu = MinkowskiReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        # same dataflow as ME's BasicBlock (conv-bn-relu-conv-bn, += residual, relu); the normalisation, the residual add
        # and the ReLU go to MinkowskiBatchNorm as ONE call each, which runs them as one pass over the rows
        residual = x
        out = self.conv1(x)
        out = self.norm1(out, act=ops.ACT_RELU)
        out = self.conv2(out)
        if self.downsample is not None:
            residual = self.downsample(x)
        if _same_map(out, residual):
            return self.norm2(out, act=ops.ACT_RELU, residual=residual.F)
        out = self.norm2(out)
        out += residual
        return self.relu(out)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, dilation=1, downsample=None, bn_momentum=0.1, dimension=-1):
        super().__init__()
        assert dimension > 0
        self.conv1 = MinkowskiConvolution(inplanes, planes, kernel_size=1, dimension=dimension)
        self.norm1 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.conv2 = MinkowskiConvolution(planes, planes, kernel_size=3, stride=stride, dilation=dilation,
                                          dimension=dimension)
        self.norm2 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.conv3 = MinkowskiConvolution(planes, planes * self.expansion, kernel_size=1, dimension=dimension)
        self.norm3 = MinkowskiBatchNorm(planes * self.expansion, momentum=bn_momentum)
        self.relu = MinkowskiReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        residual = x
        out = self.norm1(self.conv1(x), act=ops.ACT_RELU)
        out = self.norm2(self.conv2(out), act=ops.ACT_RELU)
        out = self.conv3(out)
        if self.downsample is not None:
            residual = self.downsample(x)
        if _same_map(out, residual):
            return self.norm3(out, act=ops.ACT_RELU, residual=residual.F)
        out = self.norm3(out)
        out += residual
        return self.relu(out)
