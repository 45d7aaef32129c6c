"""GPU: end-to-end class agreement with the UNMODIFIED reference over >= 1024 images (north_star: >= 99.9 % top-1).

tests/golden/eval_top1_mobilenetv2.npz: argmax / ordered top-5 of the reference over 1024 seeded images fed as four
batches of 256 in order (main_dfq.py:92-103; the QuantMeasure ranges widen from batch to batch under update_stat,
utils/quantize.py:107-111), and the reference's agreement WITH ITSELF when the input is scaled by (1 + 1e-7).
tests/golden/eval_top1_deeplab.npz: per-pixel classes of 4 x 513 x 513 images (two batches of 2).

What the goldens say about the bar: every one of the 1024 images of this random-init MobileNetV2 lands on the same
class (distinct_top1_classes = 1), so top-1 agreement is necessary but weak evidence; the ordered top-5 is the
sharper check, and there the reference agrees with its own perturbed run on only 43 % of the images (the set of
five classes is stable: 100 %).  For DeepLab the reference keeps 99.76 % of its own pixels under the perturbation,
i.e. 99.9 % is not attainable even by the reference against itself; the test asserts the measured floor minus a
margin and prints the figure."""
import json
import os

import numpy as np
import pytest
import torch

from helpers import dfq_quantised

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_mobilenetv2_top1_over_1024_images():
    from data_free_quantization_b200.eval_plan import EvalPlan
    g = np.load(os.path.join(GOLD, "eval_top1_mobilenetv2.npz"))
    info = json.load(open(os.path.join(GOLD, "eval_top1_mobilenetv2.json")))
    model, graph, bottoms, targ, tr = dfq_quantised("mobilenetv2", DEV, np.load(os.path.join(GOLD, "mobilenetv2_dfq.npz")))
    plan = EvalPlan(model, tr, batch=info["batch"])
    gen = torch.Generator().manual_seed(5)
    top1, top5, first = [], [], None
    for _ in range(info["images"] // info["batch"]):
        x = torch.randn(info["batch"], 3, 224, 224, generator=gen).to(DEV)
        logits = plan(x).view(info["batch"], -1)
        if first is None:
            first = logits[:4].cpu().numpy()
        top1.append(logits.argmax(1).cpu().numpy())
        top5.append(logits.topk(5, dim=1).indices.cpu().numpy())
    top1, top5 = np.concatenate(top1), np.concatenate(top5)
    agree = float((top1 == g["argmax"]).mean())
    set5 = float(np.mean([len(set(a) & set(b)) / 5.0 for a, b in zip(top5, g["top5"])]))
    ord5 = float((top5 == g["top5"]).all(1).mean())
    shift = float(np.abs(first - g["logits_first4"]).max() / np.abs(g["logits_first4"]).max())
    print(f"top-1 agreement {agree:.4f} over {len(top1)} images; top-5 set {set5:.4f}, ordered {ord5:.4f} "
          f"(reference vs itself under 1e-7 input noise: {info['reference_top5_set_self_agreement']:.4f} / "
          f"{info['reference_top5_ordered_self_agreement']:.4f}); logits of the first 4 images {shift:.2e} "
          f"(reference vs itself {info['reference_logit_shift_under_perturbation']:.2e})")
    assert agree >= 0.999                                   # north_star
    assert set5 >= 0.99
    # no further from the reference than the reference is from itself under a 1e-7 relative input perturbation
    assert ord5 >= info["reference_top5_ordered_self_agreement"] - 0.15
    assert shift <= 2.0 * info["reference_logit_shift_under_perturbation"]


def test_deeplab_pixel_classes_on_four_513_images():
    from data_free_quantization_b200.eval_plan import EvalPlan
    g = np.load(os.path.join(GOLD, "eval_top1_deeplab.npz"))
    info = json.load(open(os.path.join(GOLD, "eval_top1_deeplab.json")))
    model, graph, bottoms, targ, tr = dfq_quantised("deeplab", DEV, np.load(os.path.join(GOLD, "deeplab_dfq.npz")))
    plan = EvalPlan(model, tr, batch=info["batch"], in_shape=(3, 513, 513))
    gen = torch.Generator().manual_seed(5)
    am = []
    for _ in range(info["images"] // info["batch"]):
        x = torch.randn(info["batch"], 3, 513, 513, generator=gen).to(DEV)
        am.append(plan(x).view(info["batch"], 21, 513, 513).argmax(1).cpu().numpy())
    am = np.concatenate(am)
    agree = float((am == g["argmax"]).mean())
    floor = info["reference_self_agreement_under_1e-7_input_perturbation"]
    print(f"per-pixel class agreement {agree:.5f} over {am.size} pixels (reference vs itself under 1e-7 input noise: "
          f"{floor:.5f})")
    assert agree >= min(0.999, floor - 0.004)
