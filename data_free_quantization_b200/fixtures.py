"""Seeded synthetic weights for the parity tests and the bench (no checkpoints exist
offline: the reference's .pth.tar files are listed in .MISSING_LARGE_BLOBS).

Convs/linear follow the reference initialiser's distributions
(modeling/classification/MobileNetV2.py:116-130); BatchNorm statistics are randomised
 because the default gamma=1, beta=0 makes bias absorption's
c = max(0, beta - 3*gamma) identically zero.
"""
import math
from collections import OrderedDict

import torch


def synthetic_state_dict(model, seed=0, bn_seed=1):
    """Deterministic CPU state dict for `model` (same key order as model.state_dict())."""
    g = torch.Generator().manual_seed(seed)
    gb = torch.Generator().manual_seed(bn_seed)
    out = OrderedDict()
    for name, ref in model.state_dict().items():
        shape = tuple(ref.shape)
        leaf = name.rsplit(".", 1)[-1]
        if leaf == "num_batches_tracked":
            out[name] = torch.zeros((), dtype=torch.long)
        elif leaf == "running_mean":
            out[name] = torch.randn(shape, generator=gb) * 0.1
        elif leaf == "running_var":
            out[name] = torch.rand(shape, generator=gb) + 0.5
        elif len(shape) == 4:                       # conv weight, OIHW
            std = math.sqrt(2.0 / (shape[2] * shape[3] * shape[0]))
            out[name] = torch.randn(shape, generator=g) * std
        elif len(shape) == 2:                       # linear weight
            out[name] = torch.randn(shape, generator=g) * 0.01
        elif leaf == "weight":                      # BN gamma
            out[name] = torch.rand(shape, generator=gb) * 1.5 + 0.25
        elif leaf == "bias" and name.replace(".bias", ".running_mean") in model.state_dict():
            out[name] = torch.randn(shape, generator=gb) * 0.5      # BN beta
        else:                                       # conv/linear bias
            out[name] = torch.zeros(shape)
    return out
