"""DeepLabV3+ with the MobileNetV2 backbone, output stride 16 (layer shapes of the reference's segmentation model).

Like modeling/mobilenetv2.py this is a fixture, not part of the accelerated path: it supplies the 61 target-layer
shapes (38 pointwise, 17 depthwise, 6 dense k x k) and the tensor ops (residual add, F.pad, bilinear interpolate,
channel cat) of BASELINE.json configs[3].  Parameter names match the reference model
(modeling/segmentation/deeplab.py:9-34, aspp.py:32-77, decoder.py:7-45, backbone/mobilenet.py:26-118:
``backbone.features.<i>[.conv].<j>``, ``aspp.aspp<k>.atrous_conv``, ``aspp.global_avg_pool.1``, ``aspp.conv1``,
``decoder.conv1``, ``decoder.last_conv.<j>``), so a state dict moves between the two; the reference's aliasing
sub-sequences ``low_level_features`` / ``high_level_features`` are not materialised (same tensors, extra keys).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .mobilenetv2 import _STAGES


def _same_pad(x, k, dilation):
    """Explicit 'same' padding in front of an unpadded (dilated) 3x3 depthwise conv (backbone/mobilenet.py:16-22)."""
    span = k + (k - 1) * (dilation - 1) - 1
    lo = span // 2
    return F.pad(x, (lo, span - lo, lo, span - lo))


class Block(nn.Module):
    """Inverted residual whose depthwise conv carries dilation and takes its padding from F.pad on the block input."""

    def __init__(self, cin, cout, stride, dilation, t):
        super().__init__()
        hidden = round(cin * t)
        self.use_res_connect = stride == 1 and cin == cout
        self.dilation = dilation
        mods = []
        if t != 1:
            mods += [nn.Conv2d(cin, hidden, 1, bias=False), nn.BatchNorm2d(hidden), nn.ReLU6(inplace=True)]
        mods += [nn.Conv2d(hidden, hidden, 3, stride, 0, dilation, groups=hidden, bias=False),
                 nn.BatchNorm2d(hidden), nn.ReLU6(inplace=True),
                 nn.Conv2d(hidden, cout, 1, bias=False), nn.BatchNorm2d(cout)]
        self.conv = nn.Sequential(*mods)

    def forward(self, x):
        y = self.conv(_same_pad(x, 3, self.dilation))
        return x + y if self.use_res_connect else y


class Backbone(nn.Module):
    def __init__(self, output_stride=16):
        super().__init__()
        feats = [nn.Sequential(nn.Conv2d(3, 32, 3, 2, 1, bias=False), nn.BatchNorm2d(32), nn.ReLU6(inplace=True))]
        cin, reached, rate = 32, 2, 1
        for t, c, n, s in _STAGES:
            if reached == output_stride:            # keep the resolution, grow the dilation instead
                stride, dilation = 1, rate
                rate *= s
            else:
                stride, dilation = s, 1
                reached *= s
            for i in range(n):
                feats.append(Block(cin, c, stride if i == 0 else 1, dilation, t))
                cin = c
        self.features = nn.Sequential(*feats)

    def forward(self, x):
        low = None
        for i, blk in enumerate(self.features):
            x = blk(x)
            if i == 3:
                low = x                              # 24 channels at stride 4
        return x, low


class _Branch(nn.Module):
    def __init__(self, cin, cout, k, dilation):
        super().__init__()
        self.atrous_conv = nn.Conv2d(cin, cout, k, 1, 0 if k == 1 else dilation, dilation, bias=False)
        self.bn = nn.BatchNorm2d(cout)
        self.relu = nn.ReLU()

    def forward(self, x):
        return self.relu(self.bn(self.atrous_conv(x)))


class ASPP(nn.Module):
    def __init__(self, cin=320, dilations=(1, 6, 12, 18)):
        super().__init__()
        self.aspp1 = _Branch(cin, 256, 1, dilations[0])
        self.aspp2 = _Branch(cin, 256, 3, dilations[1])
        self.aspp3 = _Branch(cin, 256, 3, dilations[2])
        self.aspp4 = _Branch(cin, 256, 3, dilations[3])
        self.global_avg_pool = nn.Sequential(nn.AdaptiveAvgPool2d((1, 1)), nn.Conv2d(cin, 256, 1, bias=False),
                                             nn.BatchNorm2d(256), nn.ReLU())
        self.conv1 = nn.Conv2d(1280, 256, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(256)
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout(0.5)

    def forward(self, x):
        x1, x2, x3, x4 = self.aspp1(x), self.aspp2(x), self.aspp3(x), self.aspp4(x)
        x5 = self.global_avg_pool(x)
        x5 = F.interpolate(x5, size=x4.size()[2:], mode='bilinear', align_corners=True)
        x = torch.cat((x1, x2, x3, x4, x5), dim=1)
        return self.dropout(self.relu(self.bn1(self.conv1(x))))


class Decoder(nn.Module):
    def __init__(self, num_classes=21, low_channels=24):
        super().__init__()
        self.conv1 = nn.Conv2d(low_channels, 48, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(48)
        self.relu = nn.ReLU()
        self.last_conv = nn.Sequential(
            nn.Conv2d(304, 256, 3, 1, 1, bias=False), nn.BatchNorm2d(256), nn.ReLU(), nn.Dropout(0.5),
            nn.Conv2d(256, 256, 3, 1, 1, bias=False), nn.BatchNorm2d(256), nn.ReLU(), nn.Dropout(0.1),
            nn.Conv2d(256, num_classes, 1))

    def forward(self, x, low):
        low = self.relu(self.bn1(self.conv1(low)))
        x = F.interpolate(x, size=low.size()[2:], mode='bilinear', align_corners=True)
        return self.last_conv(torch.cat((x, low), dim=1))


class DeepLab(nn.Module):
    def __init__(self, backbone='mobilenet', output_stride=16, num_classes=21, sync_bn=False, freeze_bn=False):
        super().__init__()
        if backbone != 'mobilenet' or sync_bn:
            raise NotImplementedError("only the MobileNetV2 backbone with nn.BatchNorm2d is provided")
        self.backbone = Backbone(output_stride)
        self.aspp = ASPP()
        self.decoder = Decoder(num_classes)

    def forward(self, input):
        x, low = self.backbone(input)
        x = self.decoder(self.aspp(x), low)
        return F.interpolate(x, size=input.size()[2:], mode='bilinear', align_corners=True)
