"""Generate tests/golden/minmax_random_bn.json from the UNMODIFIED reference's set_quant_minmax.

Build container only (needs /root/reference).  Every BatchNorm gets seeded random fake_weight / fake_bias
(the quantities set_quant_minmax reads), so the 1-to-1, add and mean branches are all exercised.
Usage: python -m oracle.gen_minmax_golden
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as RH                                       # noqa: E402
from data_free_quantization_b200.fixtures import synthetic_state_dict     # noqa: E402

SEED = 7


def main():
    ns = RH.import_reference()
    torch.manual_seed(0)
    model = ns.mobilenet_v2(None)
    model.load_state_dict(synthetic_state_dict(model, seed=0, bn_seed=1))
    run = RH.RefRun(ns, model, torch.ones(4, 3, 224, 224))
    g = torch.Generator().manual_seed(SEED)
    for _name, m in run.model.named_modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.register_buffer("fake_weight", torch.rand(m.num_features, generator=g) * 1.5 + 0.25)
            m.register_buffer("fake_bias", torch.randn(m.num_features, generator=g) * 0.5)
    run.set_quant_minmax()
    ranges = {n: [float(m.running_min), float(m.running_max)] for n, m in run.model.named_modules()
              if type(m).__name__ == "QuantMeasure"}
    out = {"seed": SEED, "ranges": ranges}
    with open(os.path.join(ROOT, "tests", "golden", "minmax_random_bn.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(len(ranges), "QuantMeasure ranges written")


if __name__ == "__main__":
    main()
