"""MobileNetV2 fixture (layer shapes of the reference's classification model).

Not part of the accelerated path: it only supplies the 53 target-layer shapes the
DFQ kernels are sized on.  Parameter names match the reference model
(modeling/classification/MobileNetV2.py:68-142: ``features.<i>[.conv].<j>``,
``classifier``) so a state dict moves between the two unchanged.
"""
import torch
import torch.nn as nn

# (expansion t, out channels c, repeats n, first stride s) -- MobileNetV2.py:75-84
_STAGES = ((1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2),
           (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1))


def _cbr(cin, cout, k, stride, groups=1, act=True):
    mods = [nn.Conv2d(cin, cout, k, stride, k // 2, groups=groups, bias=False), nn.BatchNorm2d(cout)]
    if act:
        mods.append(nn.ReLU6(inplace=True))
    return mods


class InvertedResidual(nn.Module):
    def __init__(self, cin, cout, stride, t):
        super().__init__()
        hidden = cin * t
        self.use_res_connect = stride == 1 and cin == cout
        mods = [] if t == 1 else _cbr(cin, hidden, 1, 1)
        mods += _cbr(hidden, hidden, 3, stride, groups=hidden)
        mods += _cbr(hidden, cout, 1, 1, act=False)
        self.conv = nn.Sequential(*mods)

    def forward(self, x):
        if self.use_res_connect:
            return x + self.conv(x)
        return self.conv(x)


class MobileNetV2(nn.Module):
    def __init__(self, n_class=1000, last_channel=1280):
        super().__init__()
        feats = [nn.Sequential(*_cbr(3, 32, 3, 2))]
        cin = 32
        for t, c, n, s in _STAGES:
            for i in range(n):
                feats.append(InvertedResidual(cin, c, s if i == 0 else 1, t))
                cin = c
        feats.append(nn.Sequential(*_cbr(cin, last_channel, 1, 1)))
        self.features = nn.Sequential(*feats)
        self.classifier = nn.Linear(last_channel, n_class)

    def forward(self, x):
        x = self.features(x)
        x = torch.mean(x.view(x.size(0), x.size(1), -1), -1)
        return self.classifier(x)


def mobilenet_v2(path_weight=None):
    model = MobileNetV2()
    if path_weight is not None:
        model.load_state_dict(torch.load(path_weight))
    return model
